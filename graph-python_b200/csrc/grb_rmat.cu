// grb_rmat.cu — Graph500-style R-MAT generator on the device (SURVEY 8d "Synthetic inputs").
// Restates, bit for bit, the integer arithmetic of oracle/grb_oracle.c:orc_rmat_edges and
// oracle/oracle.py:edge_value, so the graph built here in HBM equals the one the oracle builds on the host
// (tests/test_gpu_rmat.py compares them).  Bench utility: nothing in the reference generates graphs.
#include <cub/cub.cuh>

#include "grb_internal.cuh"
#include "grb_storage.cuh"

#define HAVE_RMAT 1

__host__ __device__ static inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ static inline uint64_t rmat_permute(uint64_t x, int scale, uint64_t k1, uint64_t k2) {
    const uint64_t mask = (scale >= 64) ? ~0ull : ((1ull << scale) - 1);
    const int h = scale / 2 > 0 ? scale / 2 : 1;
    x = (x * k1) & mask; x ^= x >> h;
    x = (x * k2) & mask; x ^= x >> h;
    x = (x + (k1 >> 7)) & mask;
    return x;
}
__host__ __device__ static inline void rmat_edge(uint64_t e, int scale, uint64_t seed, uint64_t k1, uint64_t k2,
                                                 uint64_t *src, uint64_t *dst) {
    const uint64_t TA = 2448131358ull, TAB = 3264175144ull, TABC = 4080218931ull;
    const uint64_t st = splitmix64(seed ^ (e * 0xD1342543DE82EF95ull));
    uint64_t r = 0, c = 0;
    for (int l = 0; l < scale; l++) {
        const uint64_t u = splitmix64(st + (uint64_t)l) >> 32;
        const uint64_t rb = (u >= TAB), cb = (u >= TA && u < TAB) || (u >= TABC);
        r = (r << 1) | rb; c = (c << 1) | cb;
    }
    *src = rmat_permute(r, scale, k1, k2);
    *dst = rmat_permute(c, scale, k1, k2);
}

// keys[e] = src*n + dst (and keys[ne+e] = dst*n + src when symmetrising); self-loops become ~0 and are dropped
__global__ void k_rmat_keys(uint64_t ne, int scale, uint64_t seed, uint64_t k1, uint64_t k2, int symmetrize,
                            uint64_t *keys) {
    const uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (e >= ne) return;
    uint64_t s, d;
    rmat_edge(e, scale, seed, k1, k2, &s, &d);
    const bool loop = (s == d);
    keys[e] = loop ? ~0ull : ((s << scale) | d);
    if (symmetrize) keys[ne + e] = loop ? ~0ull : ((d << scale) | s);
}
struct NotSentinel {
    __host__ __device__ bool operator()(const uint64_t &k) const { return k != ~0ull; }
};

__host__ __device__ static inline uint64_t edge_hash(uint64_t r, uint64_t c) {
    uint64_t h = (r * 0x9E3779B97F4A7C15ull) ^ (c * 0xC2B2AE3D27D4EB4Full);
    h ^= h >> 29;
    h *= 0xBF58476D1CE4E5B9ull;
    h ^= h >> 32;
    return h;
}
template <class T>
__global__ void k_edge_values(const uint32_t *rows, const uint32_t *cols, int64_t nnz, int mode, T *vals) {
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= nnz) return;
    const uint64_t h = edge_hash(rows[p], cols[p]);
    if (mode == 1) vals[p] = (T)(1 + h % 255);
    else {
        // ((h >> 40) + 1) / 2^24 evaluated in double then rounded to T, as numpy does in oracle.edge_value
        if constexpr (std::is_same<T, float>::value) vals[p] = (float)((double)(h >> 40) + 1.0) / 16777216.0f;
        else if constexpr (std::is_same<T, double>::value) vals[p] = ((double)(h >> 40) + 1.0) / 16777216.0;
        else vals[p] = (T)(((double)(h >> 40) + 1.0) / 16777216.0);
    }
}

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

extern "C" {

GrB_Info GxB_b200_Matrix_rmat(GrB_Matrix *A, GrB_Type type, int scale, int edgefactor, uint64_t seed,
                              int symmetrize, int value_mode) {
    if (!A || !type) return GrB_NULL_POINTER;
    *A = nullptr;
    GB_Object *o = nullptr;
    API_BEGIN
    require_init();
    if (scale < 1 || scale > 31 || edgefactor < 1 || value_mode < 0 || value_mode > 2)
        grb_throw(GrB_INVALID_VALUE, "rmat: scale must be 1..31, edgefactor >= 1, value_mode 0..2");
    const int64_t n = (int64_t)1 << scale;
    const uint64_t ne = (uint64_t)edgefactor << scale;
    const int64_t nk = (int64_t)(symmetrize ? 2 * ne : ne);
    const uint64_t k1 = splitmix64(seed ^ 0xA5A5A5A5ull) | 1ull, k2 = splitmix64(seed ^ 0x5A5A5A5Aull) | 1ull;
    uint64_t *keys = dalloc<uint64_t>(nk), *kept = dalloc<uint64_t>(nk);
    GRB_LAUNCH("rmat_keys", k_rmat_keys, nblk((int64_t)ne), 256, 0, ne, scale, seed, k1, k2, symmetrize, keys);
    int64_t *d_n = g.d_scratch + 8;
    size_t tb = 0;
    CUDA_TRY(cub::DeviceSelect::If(nullptr, tb, keys, kept, d_n, nk, NotSentinel(), g.stream));
    void *tmp = dmalloc(tb);
    CUDA_TRY(cub::DeviceSelect::If(tmp, tb, keys, kept, d_n, nk, NotSentinel(), g.stream));
    g.launches += 1;
    dfree(tmp);
    const int64_t nkept = read_i64(d_n);
    dfree(keys);
    const int tcode = type->code;
    void *one = dmalloc(16);
    GRB_TYPE_SWITCH(tcode, T, { T v = (T)1; h2d(one, &v, sizeof(T)); });
    Csr c = coo_to_csr(n, n, kept, one, true, tcode, nkept, OP_SECOND, value_mode == 0);
    dfree(kept); dfree(one);
    if (value_mode != 0) {
        uint32_t *rows = csr_row_ids(c);
        GRB_TYPE_SWITCH(tcode, T,
            GRB_LAUNCH("edge_values", k_edge_values<T>, nblk(c.nnz), 256, 0, rows, c.colidx, c.nnz, value_mode,
                       (T *)c.vals));
        dfree(rows);
    }
    o = new GB_Matrix_opaque();
    o->kind = K_MATRIX; o->type = type; o->nrows = n; o->ncols = n;
    mat_set_csr(o, c);
    *A = (GrB_Matrix)o;
    API_END(o)
}

GrB_Info GxB_b200_rmat_edges_host(uint64_t *src, uint64_t *dst, int scale, uint64_t nedges, uint64_t seed) {
    if (!src || !dst) return GrB_NULL_POINTER;
    if (scale < 1 || scale > 63) return GrB_INVALID_VALUE;
    const uint64_t k1 = splitmix64(seed ^ 0xA5A5A5A5ull) | 1ull, k2 = splitmix64(seed ^ 0x5A5A5A5Aull) | 1ull;
    for (uint64_t e = 0; e < nedges; e++) rmat_edge(e, scale, seed, k1, k2, &src[e], &dst[e]);
    return GrB_SUCCESS;
}

}  // extern "C"
