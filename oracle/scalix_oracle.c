/* scalix-b200 ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C CPU restatement of the reference's algorithms for the hot path.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this; the product (scalix_b200/, include/) never does.
 *
 * Pinning: the reference has no test suite; its examples embed the only known-answer
 * checks.  tests/test_oracle.py checks this file against all of
 * them: scan of 4,000,000 ones == i+1 (/root/reference examples/inclusive_scan.cu:56-92),
 * max over idx[3] == 4,999,999 for all 27 rows (examples/reduce_last_dim.cu:55-84),
 * sum(arr3) == 4097*numelem (examples/distributed_indexing.cu:146-228) and the
 * index-generator histogram total (examples/index_generator.cu:72-93).  Float sums are
 * asserted nowhere in the reference ("parity unpinned" there); for those the oracle is an
 * fp64 sum and the tolerance is BASELINE.json's 1e-5 relative.  On the GPU box the tests
 * additionally diff against the reference's own UM build (oracle/_ref, built by
 * oracle/Makefile from the sources where they lie under /root/reference).  Outputs of that
 * build for small instances of every driver case (run on a B200 by
 * tests/golden/make_reference_golden.sh) are committed under tests/golden/ref/ and
 * tests/test_oracle.py checks this file against them on the CPU: bit-exact for integer work
 * and data movement, 1e-5 relative for the float row sums.
 *
 * The arithmetic the reference delegates to Thrust/CUB (CUDA toolkit, un-vendored; the
 * reference pins only CUDAToolkit >= 12.0, CMakeLists.txt:11; this image: CUDA 12.9 =
 * Thrust/CUB 2.8.2) is restated from its published semantics: inclusive_scan = running
 * fold in the element type; reduce = fold seeded with `init` of an associative,
 * commutative operator.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { O_I32 = 0, O_U32 = 1, O_I64 = 2, O_U64 = 3, O_F32 = 4, O_F64 = 5 };
enum { O_PLUS = 0, O_MUL = 1, O_MAX = 2, O_MIN = 3 };

/* ---- synthetic inputs ---------------------------------------------------------------- */
/* SplitMix64 keyed by (seed, linear index): the generator BASELINE.md §4.3 names */
uint64_t oracle_splitmix64(uint64_t seed, uint64_t index) {
    uint64_t z = seed + (index + 1) * 0x9E3779B97F4A7C15ull;
    z          = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z          = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
/* uniform float in [0,1): top 24 bits */
void oracle_fill_uniform_f32(float* out, uint64_t first, uint64_t n, uint64_t seed) {
    for (uint64_t i = 0; i < n; ++i)
        out[i] = (float)(oracle_splitmix64(seed, first + i) >> 40) * (1.0f / 16777216.0f);
}
/* uniform float in [-1,1) */
void oracle_fill_symmetric_f32(float* out, uint64_t first, uint64_t n, uint64_t seed) {
    for (uint64_t i = 0; i < n; ++i)
        out[i] = (float)(oracle_splitmix64(seed, first + i) >> 40) * (2.0f / 16777216.0f) - 1.0f;
}
void oracle_fill_index_u64(uint64_t* out, uint64_t first, uint64_t n, uint64_t seed,
                           uint64_t modulus) {
    for (uint64_t i = 0; i < n; ++i) out[i] = oracle_splitmix64(seed, first + i) % modulus;
}
void oracle_fill_masked_i64(int64_t* out, uint64_t first, uint64_t n, uint64_t seed,
                            uint64_t mask) {
    for (uint64_t i = 0; i < n; ++i) out[i] = (int64_t)(oracle_splitmix64(seed, first + i) & mask);
}

/* ---- index algebra: /root/reference scalix/index.cuh:53-64 and :96-108 ------------------- */
uint64_t oracle_as_linear(const uint64_t* idx, const uint64_t* shape, int rank) {
    uint64_t lin = 0, stride = 1;
    for (int i = 0; i < rank - 1; ++i) {
        lin += idx[i] * stride;
        stride *= shape[i];
    }
    return lin + idx[rank - 1] * stride;
}
void oracle_create_from_linear(uint64_t lin, const uint64_t* shape, int rank, uint64_t* idx) {
    for (int i = 0; i < rank - 1; ++i) {
        idx[i] = lin % shape[i];
        lin /= shape[i];
    }
    idx[rank - 1] = lin;
}

/* ---- partition: /root/reference scalix/array.cuh:471-508 ---------------------------------- */
/* ceil(last/ndev) columns per device; devices past the end of the array are dropped */
int oracle_partition(uint64_t last_dim, int ndev, uint64_t* start, uint64_t* len) {
    if (ndev < 1) return -1;
    if (last_dim == 0) {
        start[0] = 0;
        len[0]   = 0;
        return 1;
    }
    uint64_t split = (last_dim + (uint64_t)ndev - 1) / (uint64_t)ndev;
    int used       = 0;
    for (uint64_t s = 0; s < last_dim && used < ndev; s += split, ++used) {
        start[used] = s;
        len[used]   = last_dim - s < split ? last_dim - s : split;
    }
    return used;
}

/* ---- fill / iota: /root/reference scalix/fill.cuh:46-48, scalix/iota.cuh:46-48 -------------- */
void oracle_fill(void* out, uint64_t n, int elem_bytes, const void* value) {
    char* o = (char*)out;
    for (uint64_t i = 0; i < n; ++i) memcpy(o + i * (uint64_t)elem_bytes, value, (size_t)elem_bytes);
}
int oracle_iota(void* out, uint64_t n, int dtype, const void* start, uint64_t first) {
    switch (dtype) {
        case O_I32:
        case O_U32: {
            uint32_t s;
            memcpy(&s, start, 4);
            for (uint64_t i = 0; i < n; ++i) ((uint32_t*)out)[i] = s + (uint32_t)(first + i);
            return 0;
        }
        case O_I64:
        case O_U64: {
            uint64_t s;
            memcpy(&s, start, 8);
            for (uint64_t i = 0; i < n; ++i) ((uint64_t*)out)[i] = s + (first + i);
            return 0;
        }
        case O_F32: {
            float s;
            memcpy(&s, start, 4);
            for (uint64_t i = 0; i < n; ++i) ((float*)out)[i] = s + (float)(first + i);
            return 0;
        }
        case O_F64: {
            double s;
            memcpy(&s, start, 8);
            for (uint64_t i = 0; i < n; ++i) ((double*)out)[i] = s + (double)(first + i);
            return 0;
        }
    }
    return -1;
}

/* ---- binary operators: /root/reference scalix/algorithm/functional.cuh:38-60 ---------------- */
/* integers wrap (computed on the unsigned type); max/min follow thrust::maximum/minimum */
#define DEF_OPS(T, U, NAME)                                                        \
    static inline T op_##NAME(int op, T a, T b) {                                  \
        switch (op) {                                                              \
            case O_PLUS: return (T)((U)a + (U)b);                                  \
            case O_MUL: return (T)((U)a * (U)b);                                   \
            case O_MAX: return a < b ? b : a;                                      \
            default: return a < b ? a : b;                                         \
        }                                                                          \
    }
DEF_OPS(int32_t, uint32_t, i32)
DEF_OPS(uint32_t, uint32_t, u32)
DEF_OPS(int64_t, uint64_t, i64)
DEF_OPS(uint64_t, uint64_t, u64)
DEF_OPS(float, float, f32)
DEF_OPS(double, double, f64)

#define DISPATCH(dtype, CALL)                                                      \
    switch (dtype) {                                                               \
        case O_I32: CALL(int32_t, i32); break;                                     \
        case O_U32: CALL(uint32_t, u32); break;                                    \
        case O_I64: CALL(int64_t, i64); break;                                     \
        case O_U64: CALL(uint64_t, u64); break;                                    \
        case O_F32: CALL(float, f32); break;                                       \
        case O_F64: CALL(double, f64); break;                                      \
        default: return -1;                                                        \
    }

/* ---- inclusive scan ------------------------------------------------------------------------ */
/* single device: thrust::inclusive_scan over the flattened slice, accumulating in the element
 * type (/root/reference scalix/algorithm/inclusive_scan.cuh:87-93) */
int oracle_inclusive_scan(const void* in, void* out, uint64_t n, int dtype, int op) {
#define SCAN(T, S)                                                                 \
    {                                                                              \
        const T* a = (const T*)in;                                                 \
        T* r       = (T*)out;                                                      \
        if (n) r[0] = a[0];                                                        \
        for (uint64_t i = 1; i < n; ++i) r[i] = op_##S(op, r[i - 1], a[i]);        \
    }
    DISPATCH(dtype, SCAN)
#undef SCAN
    return 0;
}

/* the reference's multi-device algorithm, stage by stage (ibid. :75-157):
 *   1. independent scan of every partition;
 *   2. host chain: last(d) = f(last(d-1), last(d));
 *   3. fix-up on partitions d >= 1: every element EXCEPT the last becomes f(x, last(d-1)).
 * `elems` are the partition lengths in elements. */
int oracle_inclusive_scan_partitioned(const void* in, void* out, const uint64_t* elems, int nparts,
                                      int dtype, int op) {
#define PSCAN(T, S)                                                                \
    {                                                                              \
        const T* a = (const T*)in;                                                 \
        T* r       = (T*)out;                                                      \
        uint64_t off = 0;                                                          \
        for (int d = 0; d < nparts; ++d) {                                         \
            if (elems[d]) r[off] = a[off];                                         \
            for (uint64_t i = 1; i < elems[d]; ++i)                                \
                r[off + i] = op_##S(op, r[off + i - 1], a[off + i]);               \
            off += elems[d];                                                       \
        }                                                                          \
        off = elems[0];                                                            \
        for (int d = 1; d < nparts; ++d) {                                         \
            if (elems[d] == 0) continue;                                           \
            T* last = r + off + elems[d] - 1;                                      \
            T prev  = r[off - 1];                                                  \
            *last   = op_##S(op, prev, *last);                                     \
            for (uint64_t i = 0; i + 1 < elems[d]; ++i)                            \
                r[off + i] = op_##S(op, r[off + i], prev);                         \
            off += elems[d];                                                       \
        }                                                                          \
    }
    DISPATCH(dtype, PSCAN)
#undef PSCAN
    return 0;
}

/* ---- reduce_last_dim: /root/reference scalix/algorithm/reduce_last_dim.cuh:176-211 ------------- */
/* result[s] = identity; per partition: partial[s] = fold(identity, column elements of row s);
 * result[s] = f(result[s], partial[s]).  Exact for integers / max / min.  For floating-point
 * sums use oracle_reduce_last_dim_f64acc (the tolerance oracle). */
int oracle_reduce_last_dim(const void* in, uint64_t rows, const uint64_t* part_cols, int nparts,
                           int dtype, int op, const void* identity, void* out) {
#define RLD(T, S)                                                                  \
    {                                                                              \
        const T* a = (const T*)in;                                                 \
        T* r       = (T*)out;                                                      \
        T id;                                                                      \
        memcpy(&id, identity, sizeof(T));                                          \
        for (uint64_t s = 0; s < rows; ++s) r[s] = id;                             \
        uint64_t c0 = 0;                                                           \
        for (int d = 0; d < nparts; ++d) {                                         \
            for (uint64_t s = 0; s < rows; ++s) {                                  \
                T p = id;                                                          \
                for (uint64_t c = c0; c < c0 + part_cols[d]; ++c)                  \
                    p = op_##S(op, p, a[s + rows * c]);                            \
                r[s] = op_##S(op, r[s], p);                                        \
            }                                                                      \
            c0 += part_cols[d];                                                    \
        }                                                                          \
    }
    DISPATCH(dtype, RLD)
#undef RLD
    return 0;
}

/* fp64-accumulated row sums of an fp32 matrix (the "fp64 host sum" of BASELINE.json) */
void oracle_reduce_last_dim_f64acc(const float* in, uint64_t rows, uint64_t cols, double* out) {
    for (uint64_t s = 0; s < rows; ++s) out[s] = 0.0;
    for (uint64_t c = 0; c < cols; ++c) {
        const float* col = in + c * rows;
        for (uint64_t s = 0; s < rows; ++s) out[s] += (double)col[s];
    }
}

/* whole-array reduce: /root/reference scalix/algorithm/reduce.cuh:73-82 */
int oracle_reduce(const void* in, uint64_t n, int dtype, int op, const void* identity, void* out) {
#define RED(T, S)                                                                  \
    {                                                                              \
        T acc;                                                                     \
        memcpy(&acc, identity, sizeof(T));                                         \
        for (uint64_t i = 0; i < n; ++i) acc = op_##S(op, acc, ((const T*)in)[i]); \
        memcpy(out, &acc, sizeof(T));                                              \
    }
    DISPATCH(dtype, RED)
#undef RED
    return 0;
}
double oracle_sum_f32_f64acc(const float* in, uint64_t n) {
    double acc = 0.0;
    for (uint64_t i = 0; i < n; ++i) acc += (double)in[i];
    return acc;
}

/* ---- elementwise axpy through a range launch (config 3) --------------------------------------
 * y[i] = a*x[i] + y[i]; nvcc contracts this to one FMA (-fmad=true default), so the oracle
 * uses fmaf. */
void oracle_axpy_f32(float a, const float* x, float* y, uint64_t n) {
    for (uint64_t i = 0; i < n; ++i) y[i] = fmaf(a, x[i], y[i]);
}

/* ---- gather (config 4): out[i] = src[idx[i]] ---------------------------------------------------- */
void oracle_gather(const void* src, const uint64_t* idx, uint64_t n, int elem_bytes, void* out) {
    const char* s = (const char*)src;
    char* o       = (char*)out;
    for (uint64_t i = 0; i < n; ++i)
        memcpy(o + i * (uint64_t)elem_bytes, s + idx[i] * (uint64_t)elem_bytes, (size_t)elem_bytes);
}

/* ---- index-generator launch: /root/reference scalix/detail/execute_kernel.cuh:71-98 ------------
 * every partition walks the whole generator range and applies f only to indices whose last
 * coordinate falls in its slice.  Here f = "increment histogram bin" as in
 * examples/index_generator.cu:72-93.  `gen` holds n_gen md-indices of `rank` coordinates. */
void oracle_index_generator_histogram(const uint64_t* gen, uint64_t n_gen, int rank,
                                      const uint64_t* shape, const uint64_t* part_start,
                                      const uint64_t* part_len, int nparts, int32_t* hist) {
    for (int d = 0; d < nparts; ++d) {
        for (uint64_t g = 0; g < n_gen; ++g) {
            const uint64_t* idx = gen + g * (uint64_t)rank;
            uint64_t last       = idx[rank - 1];
            if (last >= part_start[d] && last < part_start[d] + part_len[d])
                hist[oracle_as_linear(idx, shape, rank)] += 1;
        }
    }
}

/* ---- examples/distributed_indexing.cu:146-217 ----------------------------------------------------
 * three alternating random-gather accumulations; indices is {64, numelem} (first index fastest).
 * Returns sum(arr3) accumulated in double; arr2/arr3 are caller-provided and updated in place. */
double oracle_distributed_indexing(const uint64_t* indices, uint64_t numelem, float* arr2,
                                   float* arr3) {
    for (uint64_t i = 0; i < numelem; ++i) arr3[i] = 1.f;
    for (uint64_t j = 0; j < numelem; ++j)
        for (uint64_t k = 0; k < 64; ++k) arr2[j] += arr3[indices[k + 64 * j]];
    for (uint64_t j = 0; j < numelem; ++j)
        for (uint64_t k = 0; k < 64; ++k) arr3[j] += arr2[indices[k + 64 * j]];
    for (uint64_t j = 0; j < numelem; ++j)
        for (uint64_t k = 0; k < 64; ++k) arr2[j] += arr3[indices[k + 64 * j]];
    double s = 0.0;
    for (uint64_t i = 0; i < numelem; ++i) s += (double)arr3[i];
    return s;
}

/* =====================================================================================
 * Host-loop baseline: the same computations as plain multithreaded C
 * loops, one pthread per requested core.  Baseline only — Scalix has no CPU path.
 * ===================================================================================== */
typedef struct {
    int tid, nthreads;
    const void* in;
    void* out;
    uint64_t rows, cols, n;
    int dtype, op;
    double* partial;   /* [nthreads][rows] for reduce, [nthreads] totals for scan */
    const uint64_t* idx;
    float a;
    pthread_barrier_t* bar;
} job_t;

static void chunk(uint64_t n, int tid, int nt, uint64_t* b, uint64_t* e) {
    uint64_t per = (n + (uint64_t)nt - 1) / (uint64_t)nt;
    *b           = per * (uint64_t)tid < n ? per * (uint64_t)tid : n;
    *e           = *b + per < n ? *b + per : n;
}

static void* rld_f32_worker(void* arg) {
    job_t* j = (job_t*)arg;
    uint64_t b, e;
    chunk(j->cols, j->tid, j->nthreads, &b, &e);
    float* acc = (float*)calloc(j->rows, sizeof(float));
    const float* a = (const float*)j->in;
    for (uint64_t c = b; c < e; ++c) {
        const float* col = a + c * j->rows;
        for (uint64_t s = 0; s < j->rows; ++s) acc[s] += col[s];
    }
    for (uint64_t s = 0; s < j->rows; ++s) j->partial[(uint64_t)j->tid * j->rows + s] = acc[s];
    free(acc);
    return 0;
}

/* per-thread partial rows in fp32, fixed-order combine */
int hostloop_reduce_last_dim_f32(const float* in, uint64_t rows, uint64_t cols, float* out,
                                 int nthreads) {
    if (nthreads < 1) nthreads = 1;
    pthread_t* th   = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    job_t* jobs     = (job_t*)calloc((size_t)nthreads, sizeof(job_t));
    double* partial = (double*)calloc((size_t)nthreads * rows, sizeof(double));
    for (int t = 0; t < nthreads; ++t) {
        jobs[t].tid = t; jobs[t].nthreads = nthreads; jobs[t].in = in;
        jobs[t].rows = rows; jobs[t].cols = cols; jobs[t].partial = partial;
        pthread_create(&th[t], 0, rld_f32_worker, &jobs[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    for (uint64_t s = 0; s < rows; ++s) {
        float acc = 0.f;
        for (int t = 0; t < nthreads; ++t) acc += (float)partial[(uint64_t)t * rows + s];
        out[s] = acc;
    }
    free(partial); free(jobs); free(th);
    return 0;
}

/* chunked two-pass parallel inclusive plus-scan on the unsigned type of the same width */
#define DEF_SCAN_WORKER(T, NAME)                                                   \
    static void* scan_worker_##NAME(void* arg) {                                   \
        job_t* j = (job_t*)arg;                                                    \
        uint64_t b, e;                                                             \
        chunk(j->n, j->tid, j->nthreads, &b, &e);                                  \
        const T* a = (const T*)j->in;                                              \
        T* r       = (T*)j->out;                                                   \
        T acc      = 0;                                                            \
        for (uint64_t i = b; i < e; ++i) { acc += a[i]; r[i] = acc; }              \
        ((T*)j->partial)[j->tid] = acc;                                            \
        pthread_barrier_wait(j->bar);                                              \
        T carry = 0;                                                               \
        for (int t = 0; t < j->tid; ++t) carry += ((T*)j->partial)[t];             \
        if (j->tid > 0) for (uint64_t i = b; i < e; ++i) r[i] += carry;            \
        return 0;                                                                  \
    }
DEF_SCAN_WORKER(uint32_t, u32)
DEF_SCAN_WORKER(uint64_t, u64)

int hostloop_inclusive_scan_plus(const void* in, void* out, uint64_t n, int elem_bytes,
                                 int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (elem_bytes != 4 && elem_bytes != 8) return -1;
    pthread_t* th   = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    job_t* jobs     = (job_t*)calloc((size_t)nthreads, sizeof(job_t));
    double* partial = (double*)calloc((size_t)nthreads, sizeof(double));
    pthread_barrier_t bar;
    pthread_barrier_init(&bar, 0, (unsigned)nthreads);
    for (int t = 0; t < nthreads; ++t) {
        jobs[t].tid = t; jobs[t].nthreads = nthreads; jobs[t].in = in; jobs[t].out = out;
        jobs[t].n = n; jobs[t].partial = partial; jobs[t].bar = &bar;
        pthread_create(&th[t], 0, elem_bytes == 4 ? scan_worker_u32 : scan_worker_u64, &jobs[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    pthread_barrier_destroy(&bar);
    free(partial); free(jobs); free(th);
    return 0;
}

static void* axpy_worker(void* arg) {
    job_t* j = (job_t*)arg;
    uint64_t b, e;
    chunk(j->n, j->tid, j->nthreads, &b, &e);
    const float* x = (const float*)j->in;
    float* y       = (float*)j->out;
    for (uint64_t i = b; i < e; ++i) y[i] = fmaf(j->a, x[i], y[i]);
    return 0;
}
int hostloop_axpy_f32(float a, const float* x, float* y, uint64_t n, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    job_t* jobs   = (job_t*)calloc((size_t)nthreads, sizeof(job_t));
    for (int t = 0; t < nthreads; ++t) {
        jobs[t].tid = t; jobs[t].nthreads = nthreads; jobs[t].in = x; jobs[t].out = y;
        jobs[t].n = n; jobs[t].a = a;
        pthread_create(&th[t], 0, axpy_worker, &jobs[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    free(jobs); free(th);
    return 0;
}

static void* gather_worker(void* arg) {
    job_t* j = (job_t*)arg;
    uint64_t b, e;
    chunk(j->n, j->tid, j->nthreads, &b, &e);
    const float* s = (const float*)j->in;
    float* o       = (float*)j->out;
    for (uint64_t i = b; i < e; ++i) o[i] = s[j->idx[i]];
    return 0;
}
int hostloop_gather_f32(const float* src, const uint64_t* idx, uint64_t n, float* out,
                        int nthreads) {
    if (nthreads < 1) nthreads = 1;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    job_t* jobs   = (job_t*)calloc((size_t)nthreads, sizeof(job_t));
    for (int t = 0; t < nthreads; ++t) {
        jobs[t].tid = t; jobs[t].nthreads = nthreads; jobs[t].in = src; jobs[t].out = out;
        jobs[t].n = n; jobs[t].idx = idx;
        pthread_create(&th[t], 0, gather_worker, &jobs[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    free(jobs); free(th);
    return 0;
}

/* parallel generators so bench inputs of several GiB are produced in seconds */
typedef struct { int tid, nt; float* out; uint64_t n, seed; } gen_t;
static void* gen_worker(void* arg) {
    gen_t* g = (gen_t*)arg;
    uint64_t b, e;
    chunk(g->n, g->tid, g->nt, &b, &e);
    oracle_fill_uniform_f32(g->out + b, b, e - b, g->seed);
    return 0;
}
void oracle_fill_uniform_f32_mt(float* out, uint64_t n, uint64_t seed, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    gen_t* g      = (gen_t*)calloc((size_t)nthreads, sizeof(gen_t));
    for (int t = 0; t < nthreads; ++t) {
        g[t].tid = t; g[t].nt = nthreads; g[t].out = out; g[t].n = n; g[t].seed = seed;
        pthread_create(&th[t], 0, gen_worker, &g[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    free(g); free(th);
}
