// selector.h -- the reference's IDSelector family (faiss/impl/IDSelector.h:19-170,
// faiss/impl/IDSelector.cpp:17-123) for the device path. A selector is evaluated ONCE per search
// over the id column of the slab into a ROW MASK: one 32-bit word per 32-row slab block, bit =
// lane = row inside the block. The scan kernels then test one bit per row instead of calling
// is_member per candidate (faiss/IndexIVFFlat.cpp:166: `if (use_sel && !sel->is_member(ids[j]))`).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <memory>
#include <vector>

namespace b200ivf {

struct Buffer;

struct IDSelector {
    virtual ~IDSelector() = default;
    // host-side membership test (faiss_IDSelector_is_member, AuxIndexStructures_c.h:54)
    virtual bool is_member(int64_t id) const = 0;
    // mask[b] bit r = is_member(ids[32*b + r]) for b in [0, nblocks); ids of pad rows are -1 and
    // their bits are never read (rows >= list length are dropped by the scan)
    virtual void eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, cudaStream_t st) const = 0;
};

struct IDSelectorRange : IDSelector { // IDSelector.h:34-58: imin <= id < imax
    int64_t imin, imax;
    IDSelectorRange(int64_t lo, int64_t hi) : imin(lo), imax(hi) {}
    bool is_member(int64_t id) const override {
        return id >= imin && id < imax;
    }
    void eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, cudaStream_t st) const override;
};

// IDSelectorBatch / IDSelectorArray (IDSelector.h:66-107): membership in a set of ids. The
// reference uses a bloom filter + unordered_set; here the set is a sorted unique array in HBM and
// a row does one binary search.
struct IDSelectorBatch : IDSelector {
    std::vector<int64_t> sorted; // host copy
    int64_t* d_sorted = nullptr;
    int device = 0;
    IDSelectorBatch(size_t n, const int64_t* indices, int device);
    ~IDSelectorBatch() override;
    bool is_member(int64_t id) const override;
    void eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, cudaStream_t st) const override;
};

// IDSelectorBitmap (IDSelector.h:110-127): bit id of a caller bitmap of n BYTES; ids outside it
// are not members (IDSelector.cpp:117-123)
struct IDSelectorBitmap : IDSelector {
    std::vector<uint8_t> bits; // host copy
    uint8_t* d_bits = nullptr;
    int device = 0;
    IDSelectorBitmap(size_t n, const uint8_t* bitmap, int device);
    ~IDSelectorBitmap() override;
    bool is_member(int64_t id) const override;
    void eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, cudaStream_t st) const override;
};

// IDSelectorNot / And / Or / XOr (IDSelector.h:130-170): operands are borrowed, as in the reference
struct IDSelectorNot : IDSelector {
    const IDSelector* sel;
    explicit IDSelectorNot(const IDSelector* s) : sel(s) {}
    bool is_member(int64_t id) const override {
        return !sel->is_member(id);
    }
    void eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, cudaStream_t st) const override;
};

struct IDSelectorBinary : IDSelector {
    enum Op { AND = 0, OR = 1, XOR = 2 };
    Op op;
    const IDSelector *lhs, *rhs;
    IDSelectorBinary(Op o, const IDSelector* l, const IDSelector* r) : op(o), lhs(l), rhs(r) {}
    bool is_member(int64_t id) const override {
        const bool a = lhs->is_member(id), b = rhs->is_member(id);
        return op == AND ? (a && b) : op == OR ? (a || b) : (a != b);
    }
    void eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, cudaStream_t st) const override;
};

struct IDSelectorAll : IDSelector { // IDSelector.h:121-127
    bool is_member(int64_t) const override {
        return true;
    }
    void eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, cudaStream_t st) const override;
};

// IVFSearchParameters (faiss/IndexIVF.h:73-88): nprobe, max_codes, sel; quantizer_params and
// inverted_list_context have no meaning for a flat quantizer over array lists
struct SearchParametersIVF {
    const IDSelector* sel = nullptr;
    size_t nprobe = 1;
    size_t max_codes = 0;
};

} // namespace b200ivf
