/* dsb_px.cuh -- the handful of data-parallel primitives the index builder (dsb_build.cu) is written in.
 *
 * Device build (nvcc, sm_100a): for_each = a grid-stride kernel around an extended __device__ lambda; sorts, scans and
 * stream compaction come from CUB (CUDA toolkit library code: the builder's plumbing, like cuBLAS for a plain GEMM);
 * every other step of the builder is a hand-written kernel body.
 *
 * TEST-ONLY host build (-DDSB_BUILD_HOSTSIM, g++): the same builder source with these primitives as plain loops, so the
 * builder's output can be compared byte for byte with the reference's `deSAMBA index` in a container that has no GPU
 * (tests/hostsim).  The product library never contains this variant.
 */
#ifndef DSB_PX_CUH
#define DSB_PX_CUH
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef DSB_BUILD_HOSTSIM
#include <algorithm>
#include <vector>
#define PX_HD inline
#define PX_LAMBDA [=]
#else
#include <cuda_runtime.h>
#include <cub/cub.cuh>
#define PX_HD __host__ __device__ __forceinline__
#define PX_LAMBDA [=] __device__
#endif

namespace px {

extern int g_launches;          /* kernels launched (device build) */
extern char g_error[512];
extern size_t g_live_bytes, g_peak_bytes;

struct Fail {};                 /* thrown on any allocation / CUDA failure; message in g_error */

inline void fail(const char *what, const char *detail)
{
    snprintf(g_error, sizeof g_error, "%s: %s", what, detail ? detail : "");
    throw Fail();
}

#ifdef DSB_BUILD_HOSTSIM
/* ------------------------------------------------------------------ host twin */
inline void *raw_alloc(size_t bytes) { void *p = malloc(bytes ? bytes : 1); if (!p) fail("malloc", "out of memory"); return p; }
inline void raw_free(void *p) { free(p); }
inline void h2d(void *d, const void *h, size_t n) { memcpy(d, h, n); }
inline void d2h(void *h, const void *d, size_t n) { memcpy(h, d, n); }
inline void d2d(void *d, const void *s, size_t n) { memcpy(d, s, n); }
inline void zero(void *d, size_t n) { memset(d, 0, n); }
inline void fill_ff(void *d, size_t n) { memset(d, 0xff, n); }
inline void sync() {}
inline size_t device_free_bytes() { return (size_t)1 << 40; }

template <class F> void for_each(uint64_t n, F f)
{
#pragma omp parallel for schedule(static)
    for (uint64_t i = 0; i < n; i++) f(i);
}
inline uint32_t atomic_or32(uint32_t *p, uint32_t v) { return __atomic_fetch_or(p, v, __ATOMIC_RELAXED); }
inline uint32_t atomic_add32(uint32_t *p, uint32_t v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
inline uint64_t atomic_add64(uint64_t *p, uint64_t v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
inline void atomic_min64(uint64_t *p, uint64_t v)
{
    uint64_t o = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (v < o && !__atomic_compare_exchange_n(p, &o, v, 1, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
}
inline void atomic_max64(uint64_t *p, uint64_t v)
{
    uint64_t o = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (v > o && !__atomic_compare_exchange_n(p, &o, v, 1, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
}

/* ascending sort of the low `bits` bits (keys never have higher bits set) */
inline void sort_keys(uint64_t *&keys, uint64_t *&alt, uint64_t n, int bits) { (void)alt; (void)bits; std::sort(keys, keys + n); }
/* stable ascending sort by key */
template <class K, class V> void sort_pairs(K *&keys, K *&kalt, V *&vals, V *&valt, uint64_t n, int bits)
{
    (void)bits;
    std::vector<uint64_t> ix(n);
    for (uint64_t i = 0; i < n; i++) ix[i] = i;
    K *k = keys;
    std::stable_sort(ix.begin(), ix.end(), [k](uint64_t a, uint64_t b) { return k[a] < k[b]; });
    for (uint64_t i = 0; i < n; i++) { kalt[i] = keys[ix[i]]; valt[i] = vals[ix[i]]; }
    std::swap(keys, kalt);
    std::swap(vals, valt);
}
/* out[i] = sum_{j<i} f(j); returns the total */
template <class T, class F> uint64_t exclusive_sum(uint64_t n, F f, T *out)
{
    uint64_t s = 0;
    for (uint64_t i = 0; i < n; i++) { uint64_t v = f(i); out[i] = (T)s; s += v; }
    return s;
}
/* out[i] = max_{j<=i} f(j) */
template <class F> void inclusive_max(uint64_t n, F f, uint64_t *out)
{
    uint64_t m = 0;
    for (uint64_t i = 0; i < n; i++) { uint64_t v = f(i); if (v > m) m = v; out[i] = m; }
}
/* out[i] = max_{j<i} in[j] (0 for i == 0); returns max over all */
inline uint64_t exclusive_max(uint64_t n, const uint64_t *in, uint64_t *out)
{
    uint64_t m = 0;
    for (uint64_t i = 0; i < n; i++) { uint64_t v = in[i]; out[i] = m; if (v > m) m = v; }
    return m;
}
/* keep the first of each run of equal values; returns the count */
inline uint64_t unique(const uint64_t *in, uint64_t n, uint64_t *out)
{
    uint64_t m = 0;
    for (uint64_t i = 0; i < n; i++) if (i == 0 || in[i] != in[i - 1]) out[m++] = in[i];
    return m;
}
/* out = { i : pred(i) } ascending; returns the count */
template <class P> uint64_t select_index(uint64_t n, P pred, uint64_t *out)
{
    uint64_t m = 0;
    for (uint64_t i = 0; i < n; i++) if (pred(i)) out[m++] = i;
    return m;
}
#else
/* ------------------------------------------------------------------ device */
#define PX_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) px::fail(#call, cudaGetErrorString(e_)); } while (0)

inline void *raw_alloc(size_t bytes)
{
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 256);
    if (e != cudaSuccess) {
        char m[160];
        snprintf(m, sizeof m, "%s (%.2f GB requested, %.2f GB live)", cudaGetErrorString(e), bytes / 1e9, g_live_bytes / 1e9);
        fail("cudaMalloc", m);
    }
    return p;
}
inline void raw_free(void *p) { if (p) cudaFree(p); }
inline void h2d(void *d, const void *h, size_t n) { PX_CUDA(cudaMemcpy(d, h, n, cudaMemcpyHostToDevice)); }
inline void d2h(void *h, const void *d, size_t n) { PX_CUDA(cudaMemcpy(h, d, n, cudaMemcpyDeviceToHost)); }
inline void d2d(void *d, const void *s, size_t n) { PX_CUDA(cudaMemcpy(d, s, n, cudaMemcpyDeviceToDevice)); }
inline void zero(void *d, size_t n) { PX_CUDA(cudaMemset(d, 0, n)); }
inline void fill_ff(void *d, size_t n) { PX_CUDA(cudaMemset(d, 0xff, n)); }
inline void sync() { PX_CUDA(cudaDeviceSynchronize()); }
inline size_t device_free_bytes() { size_t f = 0, t = 0; cudaMemGetInfo(&f, &t); return f; }

template <class F> __global__ void __launch_bounds__(256) k_build_for(uint64_t n, F f)
{
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) f(i);
}
template <class F> void for_each(uint64_t n, F f)
{
    if (!n) return;
    /* grid = a multiple of the 148 SMs x 8 resident blocks, grid-stride above that */
    uint64_t blocks = (n + 255) / 256;
    const uint64_t cap = 148ull * 8 * 4;
    if (blocks > cap) blocks = cap;
    k_build_for<<<(unsigned)blocks, 256>>>(n, f);
    g_launches++;
    PX_CUDA(cudaGetLastError());
}
__device__ __forceinline__ uint32_t atomic_or32(uint32_t *p, uint32_t v) { return atomicOr(p, v); }
__device__ __forceinline__ uint32_t atomic_add32(uint32_t *p, uint32_t v) { return atomicAdd(p, v); }
__device__ __forceinline__ uint64_t atomic_add64(uint64_t *p, uint64_t v) { return atomicAdd((unsigned long long *)p, (unsigned long long)v); }
__device__ __forceinline__ void atomic_min64(uint64_t *p, uint64_t v) { atomicMin((unsigned long long *)p, (unsigned long long)v); }
__device__ __forceinline__ void atomic_max64(uint64_t *p, uint64_t v) { atomicMax((unsigned long long *)p, (unsigned long long)v); }

struct Temp {       /* CUB temporary storage, grown on demand */
    void *p = nullptr; size_t n = 0;
    void *get(size_t bytes) { if (bytes > n) { raw_free(p); p = raw_alloc(bytes); n = bytes; } return p; }
    ~Temp() { raw_free(p); }
};

inline void sort_keys(uint64_t *&keys, uint64_t *&alt, uint64_t n, int bits)
{
    cub::DoubleBuffer<uint64_t> db(keys, alt);
    size_t tb = 0;
    PX_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tb, db, (int64_t)n, 0, bits));
    Temp t;
    PX_CUDA(cub::DeviceRadixSort::SortKeys(t.get(tb), tb, db, (int64_t)n, 0, bits));
    g_launches += (bits + 7) / 8 + 1;
    sync();
    keys = db.Current();
    alt = db.Alternate();
}
template <class K, class V> void sort_pairs(K *&keys, K *&kalt, V *&vals, V *&valt, uint64_t n, int bits)
{
    cub::DoubleBuffer<K> dk(keys, kalt);
    cub::DoubleBuffer<V> dv(vals, valt);
    size_t tb = 0;
    PX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (int64_t)n, 0, bits));
    Temp t;
    PX_CUDA(cub::DeviceRadixSort::SortPairs(t.get(tb), tb, dk, dv, (int64_t)n, 0, bits));
    g_launches += (bits + 7) / 8 + 1;
    sync();
    keys = dk.Current(); kalt = dk.Alternate();
    vals = dv.Current(); valt = dv.Alternate();
}
template <class F> struct IndexFn {       /* i -> f(i) as a random-access input iterator for CUB */
    F f;
    uint64_t base;
    using value_type = uint64_t; using difference_type = int64_t; using pointer = const uint64_t *; using reference = uint64_t;
    using iterator_category = std::random_access_iterator_tag;
    __host__ __device__ IndexFn(F f_, uint64_t b = 0) : f(f_), base(b) {}
    __device__ uint64_t operator[](int64_t i) const { return (uint64_t)f(base + (uint64_t)i); }
    __device__ uint64_t operator*() const { return (uint64_t)f(base); }
    __host__ __device__ IndexFn operator+(int64_t d) const { return IndexFn(f, base + (uint64_t)d); }
    __host__ __device__ IndexFn &operator+=(int64_t d) { base += (uint64_t)d; return *this; }
    __host__ __device__ int64_t operator-(const IndexFn &o) const { return (int64_t)(base - o.base); }
};
struct MaxOp { __host__ __device__ uint64_t operator()(uint64_t a, uint64_t b) const { return a > b ? a : b; } };

template <class T, class F> uint64_t exclusive_sum(uint64_t n, F f, T *out)
{
    if (!n) return 0;
    IndexFn<F> in(f);       /* 64-bit accumulation (the iterator's value type), narrowed on store when T is 32-bit */
    size_t tb = 0;
    PX_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int64_t)n));
    Temp t;
    PX_CUDA(cub::DeviceScan::ExclusiveSum(t.get(tb), tb, in, out, (int64_t)n));
    g_launches += 2;
    /* total = last prefix + last element */
    uint64_t *d_tot = (uint64_t *)raw_alloc(8);
    for_each(1, [=] __device__(uint64_t) { *d_tot = (uint64_t)out[n - 1] + (uint64_t)f(n - 1); });
    uint64_t tot;
    d2h(&tot, d_tot, 8);
    raw_free(d_tot);
    return tot;
}
template <class F> void inclusive_max(uint64_t n, F f, uint64_t *out)
{
    if (!n) return;
    IndexFn<F> in(f);
    size_t tb = 0;
    PX_CUDA(cub::DeviceScan::InclusiveScan(nullptr, tb, in, out, MaxOp(), (int64_t)n));
    Temp t;
    PX_CUDA(cub::DeviceScan::InclusiveScan(t.get(tb), tb, in, out, MaxOp(), (int64_t)n));
    g_launches += 2;
    sync();
}
inline uint64_t exclusive_max(uint64_t n, const uint64_t *in, uint64_t *out)
{
    if (!n) return 0;
    size_t tb = 0;
    PX_CUDA(cub::DeviceScan::ExclusiveScan(nullptr, tb, in, out, MaxOp(), (uint64_t)0, (int64_t)n));
    Temp t;
    PX_CUDA(cub::DeviceScan::ExclusiveScan(t.get(tb), tb, in, out, MaxOp(), (uint64_t)0, (int64_t)n));
    g_launches += 2;
    sync();
    uint64_t a, b;
    d2h(&a, out + (n - 1), 8);
    d2h(&b, in + (n - 1), 8);
    return a > b ? a : b;
}
inline uint64_t unique(const uint64_t *in, uint64_t n, uint64_t *out)
{
    if (!n) return 0;
    uint64_t *d_cnt = (uint64_t *)raw_alloc(8);
    size_t tb = 0;
    PX_CUDA(cub::DeviceSelect::Unique(nullptr, tb, in, out, d_cnt, (int64_t)n));
    Temp t;
    PX_CUDA(cub::DeviceSelect::Unique(t.get(tb), tb, in, out, d_cnt, (int64_t)n));
    g_launches += 2;
    uint64_t m;
    d2h(&m, d_cnt, 8);
    raw_free(d_cnt);
    return m;
}
template <class P> struct PredOfIndex { P p; __device__ bool operator()(uint64_t i) const { return p(i); } };
template <class P> uint64_t select_index(uint64_t n, P pred, uint64_t *out)
{
    if (!n) return 0;
    uint64_t *d_cnt = (uint64_t *)raw_alloc(8);
    cub::CountingInputIterator<uint64_t> it(0);
    PredOfIndex<P> po{pred};
    size_t tb = 0;
    PX_CUDA(cub::DeviceSelect::If(nullptr, tb, it, out, d_cnt, (int64_t)n, po));
    Temp t;
    PX_CUDA(cub::DeviceSelect::If(t.get(tb), tb, it, out, d_cnt, (int64_t)n, po));
    g_launches += 2;
    uint64_t m;
    d2h(&m, d_cnt, 8);
    raw_free(d_cnt);
    return m;
}
#endif

/* typed allocation with live / peak accounting */
template <class T> struct Buf {
    T *p = nullptr; uint64_t n = 0;
    Buf() {}
    explicit Buf(uint64_t n_) { alloc(n_); }
    Buf(const Buf &) = delete;
    Buf &operator=(const Buf &) = delete;
    void alloc(uint64_t n_)
    {
        release();
        n = n_;
        p = (T *)raw_alloc(n * sizeof(T));
        g_live_bytes += n * sizeof(T);
        if (g_live_bytes > g_peak_bytes) g_peak_bytes = g_live_bytes;
    }
    void release() { if (p) { raw_free(p); g_live_bytes -= n * sizeof(T); p = nullptr; n = 0; } }
    ~Buf() { release(); }
    std::vector<T> host() const { std::vector<T> h(n); if (n) d2h(h.data(), p, n * sizeof(T)); return h; }
};

}  // namespace px
#endif
