// index_io.cu -- the reference's on-disk format for the one index type of this path, so that an
// index written by the reference's write_index() loads straight onto the GPU and an index built
// here can be read back by the reference. Format restated from
//   faiss/impl/index_write.cpp:78-90    index header (d, ntotal, 2 dummies, is_trained, metric)
//   faiss/impl/index_write.cpp:396-403  IndexFlat: "IxF2" / "IxFI" + header + codes as floats
//   faiss/impl/index_write.cpp:381-389  IVF header: nlist, nprobe, quantizer, direct map
//   faiss/impl/index_write.cpp:367-379  direct map: type byte, array, (hashtable pairs)
//   faiss/impl/index_write.cpp:245-285  "ilar" ArrayInvertedLists: "full" / "sprs" size tables,
//                                       then per non-empty list its codes and its ids
//   faiss/impl/index_write.cpp:635-638  "IwFl" = IVF header + inverted lists
// and the matching readers faiss/impl/index_read.cpp:62-74,155-212,442-473,782-787.
// Everything is little-endian host data; only the list payloads go to the device (through
// add_core with the list numbers given, which keeps the stored order inside every list).
#include "index.h"

#include <algorithm>

#include <cerrno>
#include <cstdio>
#include <cstring>
#include <memory>

namespace b200ivf {

namespace {

constexpr uint32_t fourcc(const char (&s)[5]) { // faiss/impl/io.cpp:237-241
    return (uint32_t)(unsigned char)s[0] | (uint32_t)(unsigned char)s[1] << 8 |
            (uint32_t)(unsigned char)s[2] << 16 | (uint32_t)(unsigned char)s[3] << 24;
}

std::string fourcc_printable(uint32_t x) {
    std::string r;
    for (int i = 0; i < 4; ++i) {
        const char c = (char)((x >> (8 * i)) & 0xff);
        r += (c >= 32 && c < 127) ? c : '?';
    }
    return r;
}

struct File {
    FILE* f = nullptr;
    std::string name;
    File(const char* fname, const char* mode) : name(fname ? fname : "") {
        if (!fname)
            throw B200Exception("null file name");
        f = std::fopen(fname, mode);
        if (!f) // faiss/impl/io.cpp FileIOReader / FileIOWriter
            throw B200Exception("could not open " + name + " for " +
                                (mode[0] == 'r' ? "reading" : "writing") + ": " +
                                std::strerror(errno));
    }
    ~File() {
        if (f)
            std::fclose(f);
    }
    void close_checked() {
        const int r = std::fclose(f);
        f = nullptr;
        if (r != 0)
            throw B200Exception("file " + name + " close error: " + std::strerror(errno));
    }
    template <class T>
    void read(T* p, size_t n) { // READANDCHECK, faiss/impl/io_macros.h:18-28
        const size_t ret = n ? std::fread(p, sizeof(T), n, f) : 0;
        if (ret != n)
            throw B200Exception("read error in " + name + ": " + std::to_string(ret) +
                                " != " + std::to_string(n) + " (" + std::strerror(errno) + ")");
    }
    template <class T>
    void write(const T* p, size_t n) { // WRITEANDCHECK, faiss/impl/io_macros.h:48-58
        const size_t ret = n ? std::fwrite(p, sizeof(T), n, f) : 0;
        if (ret != n)
            throw B200Exception("write error in " + name + ": " + std::to_string(ret) +
                                " != " + std::to_string(n) + " (" + std::strerror(errno) + ")");
    }
    template <class T>
    T read1() {
        T v;
        read(&v, 1);
        return v;
    }
    template <class T>
    void write1(const T& v) {
        write(&v, 1);
    }
    template <class T>
    void read_vector(std::vector<T>& v) { // READVECTOR, faiss/impl/io_macros.h:39-46
        const size_t size = read1<size_t>();
        if (!(size < ((uint64_t)1 << 40)))
            throw B200Exception("Error: 'size >= 0 && size < (uint64_t{1} << 40)' failed");
        v.resize(size);
        read(v.data(), size);
    }
    template <class T>
    void write_vector(const std::vector<T>& v) {
        write1<size_t>(v.size());
        write(v.data(), v.size());
    }
};

struct Header { // faiss/impl/index_write.cpp:78-90
    int d = 0;
    int64_t ntotal = 0;
    bool is_trained = false;
    int metric_type = 1;
};

Header read_header(File& f) {
    Header h;
    h.d = f.read1<int>();
    h.ntotal = f.read1<int64_t>();
    f.read1<int64_t>();
    f.read1<int64_t>();
    h.is_trained = f.read1<char>() != 0;
    h.metric_type = f.read1<int>();
    if (h.metric_type > 1)
        throw B200Exception("metric type " + std::to_string(h.metric_type) +
                            " not supported on the B200 path (METRIC_L2 / METRIC_INNER_PRODUCT)");
    return h;
}

void write_header(File& f, int d, int64_t ntotal, bool is_trained, int metric_type) {
    f.write1<int>(d);
    f.write1<int64_t>(ntotal);
    const int64_t dummy = 1 << 20;
    f.write1(dummy);
    f.write1(dummy);
    f.write1<char>(is_trained ? 1 : 0);
    f.write1<int>(metric_type);
}

} // namespace

IndexIVFFlatB200* read_index(const char* fname, int /*io_flags*/, int device) {
    File f(fname, "rb");
    const uint32_t h = f.read1<uint32_t>();
    if (h != fourcc("IwFl"))
        throw B200Exception("Index type 0x" + [&] {
            char b[16];
            std::snprintf(b, sizeof(b), "%08x", h);
            return std::string(b);
        }() + " (\"" + fourcc_printable(h) +
                            "\") not recognized: the B200 path reads IndexIVFFlat (\"IwFl\") only");
    const Header ih = read_header(f);
    const size_t nlist = f.read1<size_t>();
    const size_t nprobe = f.read1<size_t>();
    // coarse quantizer: a flat index (faiss/impl/index_read.cpp:593-611)
    const uint32_t qh = f.read1<uint32_t>();
    if (qh != fourcc("IxF2") && qh != fourcc("IxFI") && qh != fourcc("IxFl"))
        throw B200Exception("coarse quantizer type \"" + fourcc_printable(qh) +
                            "\" not supported on the B200 path (IndexFlat only)");
    const Header qhd = read_header(f);
    std::vector<float> centroids;
    f.read_vector(centroids); // READXBVECTOR: the size is in floats
    if (qhd.d != ih.d || (size_t)qhd.ntotal != nlist ||
        centroids.size() != nlist * (size_t)ih.d)
        throw B200Exception("coarse quantizer does not hold nlist x d centroids");
    if (qhd.metric_type != ih.metric_type)
        throw B200Exception("coarse quantizer metric differs from the index metric");
    // direct map (not used by search): read and drop
    const char dm_type = f.read1<char>();
    {
        std::vector<int64_t> arr;
        f.read_vector(arr);
        if (dm_type == 2) { // DirectMap::Hashtable
            const size_t npairs = f.read1<size_t>();
            if (!(npairs < ((uint64_t)1 << 40))) // the bound read_vector puts on every length field
                throw B200Exception("Error: 'npairs < (uint64_t{1} << 40)' failed");
            // dropped in bounded pieces: a corrupt count fails in read(), not in an allocation
            std::vector<int64_t> piece((size_t)1 << 16);
            for (size_t left = 2 * npairs; left > 0;) {
                const size_t take = std::min(left, piece.size());
                f.read(piece.data(), take);
                left -= take;
            }
        }
    }
    // inverted lists
    const uint32_t ih2 = f.read1<uint32_t>();
    std::unique_ptr<IndexIVFFlatB200> idx(
            new IndexIVFFlatB200((size_t)ih.d, nlist, ih.metric_type, device));
    idx->nprobe = nprobe;
    if (ih.is_trained)
        idx->set_centroids(centroids.data());
    if (ih2 == fourcc("il00")) { // no lists stored
        idx->ntotal = ih.ntotal;
        return idx.release();
    }
    if (ih2 != fourcc("ilar"))
        throw B200Exception("inverted list type \"" + fourcc_printable(ih2) +
                            "\" not supported on the B200 path (ArrayInvertedLists only)");
    const size_t il_nlist = f.read1<size_t>();
    const size_t code_size = f.read1<size_t>();
    if (il_nlist != nlist) // faiss/impl/index_read.cpp:234
        throw B200Exception("Error: 'ils->nlist == ivf->nlist' failed");
    if (code_size != (size_t)ih.d * sizeof(float))
        throw B200Exception("Error: 'ils->code_size == ivf->code_size' failed");
    std::vector<size_t> sizes(nlist, 0);
    { // read_ArrayInvertedLists_sizes, faiss/impl/index_read.cpp:155-183
        const uint32_t list_type = f.read1<uint32_t>();
        std::vector<size_t> tab;
        f.read_vector(tab);
        if (list_type == fourcc("full")) {
            if (tab.size() != nlist)
                throw B200Exception("Error: 'sizes.size() == nlist' failed");
            sizes = tab;
        } else if (list_type == fourcc("sprs")) {
            if (tab.size() % 2 != 0)
                throw B200Exception("Error: 'idsizes.size() % 2 == 0' failed");
            for (size_t j = 0; j < tab.size(); j += 2) {
                if (tab[j] >= nlist)
                    throw B200Exception("Error: 'idsizes[j] < sizes.size()' failed");
                sizes[tab[j]] = tab[j + 1];
            }
        } else {
            throw B200Exception("list_type \"" + fourcc_printable(list_type) + "\" not recognized");
        }
    }
    if (!ih.is_trained)
        throw B200Exception("index file holds inverted lists but is not trained");
    // payload: groups of whole lists of about kGroup vectors each go through add_core with their
    // list numbers; the stored order inside a list is kept
    const size_t kGroup = (size_t)1 << 20;
    const size_t d = (size_t)ih.d;
    {
        // the sizes are known up front: one slab, filled in place
        std::vector<int64_t> rs(nlist);
        for (size_t j = 0; j < nlist; ++j)
            rs[j] = (int64_t)sizes[j];
        idx->reserve_lists(rs.data());
    }
    std::vector<float> codes;
    std::vector<int64_t> ids, assign;
    size_t l = 0;
    int64_t total = 0;
    while (l < nlist) {
        size_t l1 = l, cnt = 0;
        while (l1 < nlist && (cnt == 0 || cnt + sizes[l1] <= kGroup))
            cnt += sizes[l1++];
        if (cnt > 0) {
            codes.resize(cnt * d);
            ids.resize(cnt);
            assign.resize(cnt);
            size_t o = 0;
            for (size_t j = l; j < l1; ++j) {
                const size_t n = sizes[j];
                if (n == 0)
                    continue;
                f.read(codes.data() + o * d, n * d);
                f.read(ids.data() + o, n);
                std::fill(assign.begin() + o, assign.begin() + o + n, (int64_t)j);
                o += n;
            }
            idx->add_core((int64_t)cnt, codes.data(), ids.data(), assign.data());
            total += (int64_t)cnt;
        }
        l = l1;
    }
    idx->ntotal = ih.ntotal; // the reference counts skipped (list -1) vectors too
    (void)total;
    return idx.release();
}

void write_index(const IndexIVFFlatB200* idx, const char* fname) {
    if (!idx)
        throw B200Exception("null index");
    File f(fname, "wb");
    const size_t nlist = idx->nlist, d = (size_t)idx->d;
    f.write1<uint32_t>(fourcc("IwFl"));
    write_header(f, idx->d, idx->ntotal, idx->is_trained, idx->metric_type);
    f.write1<size_t>(nlist);
    f.write1<size_t>(idx->nprobe);
    // quantizer
    f.write1<uint32_t>(idx->metric_type == 0 /* METRIC_INNER_PRODUCT */ ? fourcc("IxFI") : fourcc("IxF2"));
    write_header(f, idx->d, idx->is_trained ? (int64_t)nlist : 0, true, idx->metric_type);
    std::vector<float> centroids;
    if (idx->is_trained) {
        centroids.resize(nlist * d);
        idx->get_centroids(centroids.data());
    }
    f.write_vector(centroids);
    // direct map: DirectMap::NoMap, empty array
    f.write1<char>(0);
    f.write1<size_t>(0);
    // inverted lists
    f.write1<uint32_t>(fourcc("ilar"));
    f.write1<size_t>(nlist);
    f.write1<size_t>(d * sizeof(float));
    std::vector<size_t> sizes(nlist);
    size_t n_non0 = 0;
    for (size_t l = 0; l < nlist; ++l) {
        sizes[l] = idx->get_list_size(l);
        n_non0 += sizes[l] > 0;
    }
    if (n_non0 > nlist / 2) {
        f.write1<uint32_t>(fourcc("full"));
        f.write_vector(sizes);
    } else {
        f.write1<uint32_t>(fourcc("sprs"));
        std::vector<size_t> tab;
        for (size_t l = 0; l < nlist; ++l)
            if (sizes[l] > 0) {
                tab.push_back(l);
                tab.push_back(sizes[l]);
            }
        f.write_vector(tab);
    }
    std::vector<float> codes;
    std::vector<int64_t> ids;
    for (size_t l = 0; l < nlist; ++l) {
        const size_t n = sizes[l];
        if (n == 0)
            continue;
        codes.resize(n * d);
        ids.resize(n);
        idx->get_list_codes(l, codes.data());
        idx->get_list_ids(l, ids.data());
        f.write(codes.data(), n * d);
        f.write(ids.data(), n);
    }
    f.close_checked();
}

} // namespace b200ivf
