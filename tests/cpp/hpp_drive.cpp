// tests/cpp/hpp_drive.cpp -- drives include/b200ivf.hpp the way the fork's benchmark drives
// faiss::IndexIVFFlat (tutorial/cpp/benchmark-ivf/benchmark_ivf.cpp:250-320: build, add, set
// nprobe, index.search(nq, xq, k, D, I)). Usage: hpp_drive in.bin out.bin
//   in.bin : int64 {d, nb, nlist, nq, k, nprobe, metric}, float radius, then centroids, xb, xq
//   out.bin: D, I (plain search) | D, I (IDSelectorRange(0, nb/2)) | lims, range labels, distances
// Without a CUDA device the constructor throws; the program then prints NO_DEVICE and exits 0.
#include "b200ivf.hpp"

#include <cstdio>
#include <vector>

int main(int argc, char** argv) {
    if (argc < 3)
        return 2;
    FILE* f = fopen(argv[1], "rb");
    if (!f)
        return 2;
    int64_t hd[7];
    float radius;
    if (fread(hd, sizeof(int64_t), 7, f) != 7 || fread(&radius, sizeof(float), 1, f) != 1)
        return 2;
    const int64_t d = hd[0], nb = hd[1], nlist = hd[2], nq = hd[3], k = hd[4], nprobe = hd[5];
    std::vector<float> cent(nlist * d), xb(nb * d), xq(nq * d);
    if (fread(cent.data(), 4, cent.size(), f) != cent.size() ||
        fread(xb.data(), 4, xb.size(), f) != xb.size() || fread(xq.data(), 4, xq.size(), f) != xq.size())
        return 2;
    fclose(f);
    try {
        b200::IndexIVFFlat index(d, nlist, (b200::MetricType)hd[6], 0);
        index.set_centroids(cent.data());
        index.add(nb, xb.data());
        index.nprobe = nprobe;
        std::vector<float> D(nq * k), Ds(nq * k);
        std::vector<b200::idx_t> I(nq * k), Is(nq * k);
        index.search(nq, xq.data(), k, D.data(), I.data());
        b200::IDSelectorRange sel(0, nb / 2);
        b200::SearchParametersIVF p;
        p.nprobe = nprobe;
        p.sel = &sel;
        index.search(nq, xq.data(), k, Ds.data(), Is.data(), &p);
        b200::RangeSearchResult res(nq);
        index.range_search(nq, xq.data(), radius, &res);
        FILE* o = fopen(argv[2], "wb");
        fwrite(D.data(), 4, D.size(), o);
        fwrite(I.data(), 8, I.size(), o);
        fwrite(Ds.data(), 4, Ds.size(), o);
        fwrite(Is.data(), 8, Is.size(), o);
        fwrite(res.lims.data(), sizeof(size_t), res.lims.size(), o);
        fwrite(res.labels.data(), 8, res.labels.size(), o);
        fwrite(res.distances.data(), 4, res.distances.size(), o);
        fclose(o);
        printf("OK ntotal=%lld trained=%d results=%zu\n", (long long)index.ntotal(),
               (int)index.is_trained(), res.labels.size());
        // the sharded face with a world of one rank (no NCCL needed): same vectors, same result
        {
            b200::IndexShardsIVFFlat sh(d, nlist, (b200::MetricType)hd[6], 0, 0, 1, nullptr);
            sh.set_centroids(cent.data());
            std::vector<b200::idx_t> ids(nb);
            for (int64_t i = 0; i < nb; ++i)
                ids[i] = i;
            sh.add_with_ids(nb, xb.data(), ids.data());
            sh.nprobe = nprobe;
            std::vector<float> D2(nq * k);
            std::vector<b200::idx_t> I2(nq * k);
            sh.search(nq, xq.data(), k, D2.data(), I2.data());
            sh.synchronize();
            if (I2 != I || D2 != D || sh.ntotal() != nb || sh.world() != 1) {
                printf("SHARDED_MISMATCH\n");
                return 1;
            }
            printf("SHARDED_OK\n");
        }
        // error path: k = 0 must throw with the reference's message shape (faiss/IndexIVF.cpp:310)
        try {
            index.search(nq, xq.data(), 0, D.data(), I.data());
            printf("NO_THROW\n");
            return 1;
        } catch (const b200::Exception& e) {
            printf("THROWS code=%d %s\n", e.code, e.what());
        }
    } catch (const b200::Exception& e) {
        if (std::string(e.what()).find("no CUDA device") != std::string::npos) {
            printf("NO_DEVICE code=%d\n", e.code);
            return 0;
        }
        printf("ERROR %d %s\n", e.code, e.what());
        return 1;
    }
    return 0;
}
