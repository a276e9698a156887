// pool.cu - fused masked mean-pool + L2 normalise of encoder hidden states.
//
// Stands behind assistant/ai/embedders/transformers.py:25 of the reference
// (`outputs.last_hidden_state.mean(dim=1)`), generalised to a padded batch (masked mean)
// and followed by the L2 normalisation the cosine consumer makes harmless.
//
// One pass over [B,T,D]: HBM-bound streaming, 128-bit loads, fp32 accumulation.
// Grid = (S, B) with a thread-block cluster of S CTAs along x: the S CTAs of a cluster
// split the token axis of one text, reduce through distributed shared memory, and the
// cluster's rank-0 CTA normalises and writes the row.  S is picked so that B*S covers
// the 148 SMs a few times even for small batches.
#include <cooperative_groups.h>
#include <stdlib.h>

#include "store.cuh"

namespace cg = cooperative_groups;

namespace dab {

namespace {

constexpr int kPoolThreads = 256;

__device__ __forceinline__ bool mask_on(const void *mask, int mask_dtype, long long i) {
    if (mask == nullptr) return true;
    switch (mask_dtype) {
        case DAB_MASK_U8: return static_cast<const uint8_t *>(mask)[i] != 0;
        case DAB_MASK_I32: return static_cast<const int32_t *>(mask)[i] != 0;
        default: return static_cast<const long long *>(mask)[i] != 0;
    }
}

__device__ __forceinline__ uint4 ld_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

template <bool BF16>
struct Vec;
template <>
struct Vec<false> {  // 4 fp32 per 16 bytes
    static constexpr int kElems = 4;
    __device__ static __forceinline__ void add(float *acc, const uint4 &v) {
        acc[0] += __uint_as_float(v.x);
        acc[1] += __uint_as_float(v.y);
        acc[2] += __uint_as_float(v.z);
        acc[3] += __uint_as_float(v.w);
    }
};
template <>
struct Vec<true> {  // 8 bf16 per 16 bytes
    static constexpr int kElems = 8;
    __device__ static __forceinline__ void add(float *acc, const uint4 &v) {
        acc[0] += bf16_lo(v.x);
        acc[1] += bf16_hi(v.x);
        acc[2] += bf16_lo(v.y);
        acc[3] += bf16_hi(v.y);
        acc[4] += bf16_lo(v.z);
        acc[5] += bf16_hi(v.z);
        acc[6] += bf16_lo(v.w);
        acc[7] += bf16_hi(v.w);
    }
};

constexpr int kTokChunk = 2048;  // tokens whose mask is compacted into shared memory at a time

// NV = 16-byte column vectors per thread (1 when D*elem <= 4 KiB, up to 4).
//
// Per text: the mask is compacted, chunk by chunk, into a shared list of the token rows that count;
// the S CTAs of the cluster and the row groups inside a CTA then take entries of that list round
// robin.  Masked-out rows cost nothing (never loaded), work is balanced however ragged the batch
// is, and no load waits on a mask load.
// Measured on the C5 batch (256 x 512 x 1024 fp32, lengths 16..512): cluster sizes 2 / 4 / 8 give 55.9 / 54.5 / 56.8 us,
// and 4 / 12 / 16 rows in flight per thread 57.8 / 56.2 / 63.6 us against 54.4 us for this form (8 rows, 4 CTAs per SM):
// the step is bounded by the ~13 us every launch of this size pays for ramp-up and tail (dense batches: 86 us at
// B = 256, 305 us at B = 1024, i.e. 7.4 TB/s marginal), not by the access pattern.
template <bool BF16, int NV>
__global__ void __launch_bounds__(kPoolThreads)
pool_l2norm_kernel(const uint4 *__restrict__ hidden, const void *__restrict__ mask, int mask_dtype,
                   float *__restrict__ out, __nv_bfloat16 *__restrict__ out_bf16, float *__restrict__ out_inv,
                   int T, int D, int vecs_per_row, int row_groups, StoreSink sink) {
    constexpr int VE = Vec<BF16>::kElems;
    constexpr int U = 8 / NV;  // rows in flight per thread
    extern __shared__ float smem[];
    float *part = smem;                                    // [row_groups][D] partial sums of this CTA
    float *psum = smem + (size_t)row_groups * D;           // [D] CTA total, read by rank 0 over DSMEM
    __shared__ unsigned short s_rows[kTokChunk];           // compacted offsets (within the chunk) of counted tokens
    __shared__ int s_warp_tot[kPoolThreads / 32];
    __shared__ float s_red[kPoolThreads / 32];
    __shared__ float s_inv;

    cg::cluster_group cluster = cg::this_cluster();
    const int S = gridDim.x;
    const int rank = blockIdx.x;
    const long long b = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rg = NV == 1 ? tid / vecs_per_row : 0;       // NV > 1 => one row group
    const int c0 = NV == 1 ? tid % vecs_per_row : tid;
    const bool active = NV == 1 ? (tid < row_groups * vecs_per_row) : true;
    const int lanes_total = S * row_groups;                // list entries are dealt to (rank, row group) pairs
    const int my_lane = rank * row_groups + rg;

    float acc[NV][VE];
#pragma unroll
    for (int v = 0; v < NV; v++)
#pragma unroll
        for (int e = 0; e < VE; e++) acc[v][e] = 0.f;

    int total = 0;  // counted tokens of the whole text (every CTA of the cluster computes the same number)
    const uint4 *base = hidden + (size_t)b * T * vecs_per_row;
    for (int chunk0 = 0; chunk0 < T; chunk0 += kTokChunk) {
        const int clen = min(kTokChunk, T - chunk0);
        int n_c;
        if (mask == nullptr) {
            n_c = clen;
        } else {
            // ordered compaction of the chunk's mask: ballot + warp prefix + per-warp totals
            n_c = 0;
            for (int t0 = 0; t0 < clen; t0 += kPoolThreads) {
                const int t = t0 + tid;
                const bool on = t < clen && mask_on(mask, mask_dtype, b * T + chunk0 + t);
                const unsigned bal = __ballot_sync(0xffffffffu, on);
                if (lane == 0) s_warp_tot[warp] = __popc(bal);
                __syncthreads();
                int off = n_c, blk = 0;
#pragma unroll
                for (int w = 0; w < kPoolThreads / 32; w++) {
                    const int c = s_warp_tot[w];
                    if (w < warp) off += c;
                    blk += c;
                }
                if (on) s_rows[off + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)t;
                n_c += blk;
                __syncthreads();
            }
        }
        total += n_c;
        if (active) {
            for (int j = my_lane; j < n_c; j += lanes_total * U) {
                uint4 vals[U][NV];
                bool on[U];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int jj = j + u * lanes_total;
                    on[u] = jj < n_c;
                    if (on[u]) {
                        const int t = chunk0 + (mask == nullptr ? jj : (int)s_rows[jj]);
                        const uint4 *row = base + (size_t)t * vecs_per_row;
#pragma unroll
                        for (int v = 0; v < NV; v++) {
                            const int c = c0 + v * kPoolThreads;
                            if (NV == 1 || c < vecs_per_row) vals[u][v] = ld_stream(row + c);
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    if (on[u]) {
#pragma unroll
                        for (int v = 0; v < NV; v++) {
                            const int c = c0 + v * kPoolThreads;
                            if (NV == 1 || c < vecs_per_row) Vec<BF16>::add(acc[v], vals[u][v]);
                        }
                    }
                }
            }
        }
        __syncthreads();  // s_rows is rewritten by the next chunk
    }
    if (active) {
#pragma unroll
        for (int v = 0; v < NV; v++) {
            const int c = c0 + v * kPoolThreads;
            if (NV == 1 || c < vecs_per_row) {
#pragma unroll
                for (int e = 0; e < VE; e++) part[(size_t)rg * D + (size_t)c * VE + e] = acc[v][e];
            }
        }
    }
    __syncthreads();
    // fixed-order reduction over the row groups -> psum
    for (int d = tid; d < D; d += kPoolThreads) {
        float s = 0.f;
        for (int g = 0; g < row_groups; g++) s += part[(size_t)g * D + d];
        psum[d] = s;
    }
    cluster.sync();  // every CTA's psum is visible cluster-wide

    if (rank == 0) {
        const float inv_cnt = 1.0f / (float)max(total, 1);
        float ss = 0.f;
        for (int d = tid; d < D; d += kPoolThreads) {
            float s = 0.f;
            for (int r = 0; r < S; r++) s += cluster.map_shared_rank(psum, r)[d];
            s *= inv_cnt;
            part[d] = s;  // reuse as the mean
            ss += s * s;
        }
        ss = warp_sum(ss);
        if (lane == 0) s_red[warp] = ss;
        __syncthreads();
        if (tid == 0) {
            float t = 0.f;
            for (int w = 0; w < kPoolThreads / 32; w++) t += s_red[w];
            s_inv = 1.0f / fmaxf(sqrtf(t), 1e-12f);
        }
        __syncthreads();
        const float inv = s_inv;
        float stored_ss = 0.f;     // |row|^2 of the values as STORED (ingest: what dab_store_append would compute)
        for (int d = tid; d < D; d += kPoolThreads) {
            const float o = part[d] * inv;
            if (out) out[(size_t)b * D + d] = o;
            float kept = o;
            if (out_bf16) {
                const __nv_bfloat16 ob = __float2bfloat16_rn(o);
                out_bf16[(size_t)b * D + d] = ob;
                kept = __bfloat162float(ob);
            }
            stored_ss = fmaf(kept, kept, stored_ss);
        }
        if (tid == 0 && out_inv) out_inv[b] = inv;
        if (sink.inv_norm != nullptr) {
            // the row lands in a store: its 1/|row| (from the stored values, 0 for a zero row) and its id go with it
            __syncthreads();               // s_red is reused
            stored_ss = warp_sum(stored_ss);
            if (lane == 0) s_red[warp] = stored_ss;
            __syncthreads();
            if (tid == 0) {
                float t = 0.f;
                for (int w = 0; w < kPoolThreads / 32; w++) t += s_red[w];
                sink.inv_norm[b] = (t > 0.f && t < INFINITY) ? 1.0f / sqrtf(t) : 0.f;
                sink.ids[b] = sink.id_first + b;
            }
        }
    }
    cluster.sync();  // keep remote shared memory alive until rank 0 has read it
}

template <bool BF16, int NV>
int launch_pool(const void *hidden, const void *mask, int mask_dtype, float *out, void *out_bf16, float *out_inv,
                int64_t B, int64_t T, int64_t D, int S, int vecs_per_row, int row_groups, cudaStream_t stream, StoreSink sink) {
    auto kern = pool_l2norm_kernel<BF16, NV>;
    const size_t smem = ((size_t)row_groups * D + D) * sizeof(float);
    if (smem > 48 * 1024) DAB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(S, (unsigned)B, 1);
    cfg.blockDim = dim3(kPoolThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = S;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DAB_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, static_cast<const uint4 *>(hidden), mask, mask_dtype, out,
                                   static_cast<__nv_bfloat16 *>(out_bf16), out_inv, (int)T, (int)D, vecs_per_row,
                                   row_groups, sink));
    count_launch();
    return DAB_OK;
}

}  // namespace

// chosen so B*S gives every SM several CTAs, never dealing fewer than ~16 tokens to a row group
int pool_pick_splits(int64_t B, int64_t T, int row_groups) {
    if (const char *e = tuning_env("DAB_POOL_SPLITS")) {  // tuning aid: force the cluster size (1, 2, 4, 8)
        const int v = atoi(e);
        if (v == 1 || v == 2 || v == 4 || v == 8) return v;
    }
    int S = 1;
    while (S < 8 && B * S < 4 * 148 && T / (2 * S) >= 16 * row_groups) S *= 2;
    return S;
}

}  // namespace dab

namespace dab {
// the pool launch behind dab_pool_l2norm and dab_store_ingest (`sink`: where the rows' inverse norms and ids go when
// the outputs are a store's own memory; `out` may then be NULL)
static int pool_run(const void *hidden, int hidden_dtype, const void *mask, int mask_dtype, float *out, void *out_bf16,
                    float *out_inv, int64_t B, int64_t T, int64_t D, void *stream, StoreSink sink) {
    if (B == 0 && T >= 0 && D > 0) return DAB_OK;
    if ((!out && !sink.inv_norm) || (!hidden && T > 0)) return fail(DAB_E_INVALID, "dab_pool_l2norm: hidden/out must not be NULL");
    if (hidden_dtype != DAB_F32 && hidden_dtype != DAB_BF16) return fail(DAB_E_INVALID, "dab_pool_l2norm: hidden_dtype must be DAB_F32 or DAB_BF16");
    if (mask && (mask_dtype < DAB_MASK_U8 || mask_dtype > DAB_MASK_I64)) return fail(DAB_E_INVALID, "dab_pool_l2norm: bad mask_dtype");
    if (B < 0 || T < 0 || D <= 0) return fail(DAB_E_INVALID, "dab_pool_l2norm: negative size");
    if (B > 65535) return fail(DAB_E_INVALID, "dab_pool_l2norm: B > 65535, split the batch");
    if (T > (1 << 24)) return fail(DAB_E_INVALID, "dab_pool_l2norm: T too large");
    const int ve = hidden_dtype == DAB_BF16 ? 8 : 4;
    if (D % ve) return fail(DAB_E_INVALID, "dab_pool_l2norm: D must be a multiple of 4 (fp32) / 8 (bf16)");
    if (!aligned16(hidden) || (out && !aligned16(out))) return fail(DAB_E_INVALID, "dab_pool_l2norm: pointers must be 16-byte aligned");
    const int vecs = (int)(D / ve);
    int nv = 1, row_groups = 1;
    if (vecs <= kPoolThreads) {
        row_groups = kPoolThreads / vecs;
    } else {
        nv = (vecs + kPoolThreads - 1) / kPoolThreads;
        if (nv > 4) return fail(DAB_E_INVALID, "dab_pool_l2norm: D too large (max 4096 fp32 / 8192 bf16)");
        if (nv == 3) nv = 4;
    }
    const int S = pool_pick_splits(B, T, row_groups);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool bf = hidden_dtype == DAB_BF16;
#define DAB_POOL_CASE(BF, NV_) return launch_pool<BF, NV_>(hidden, mask, mask_dtype, out, out_bf16, out_inv, B, T, D, S, vecs, row_groups, st, sink)
    if (nv == 1) { if (bf) DAB_POOL_CASE(true, 1); else DAB_POOL_CASE(false, 1); }
    if (nv == 2) { if (bf) DAB_POOL_CASE(true, 2); else DAB_POOL_CASE(false, 2); }
    if (bf) DAB_POOL_CASE(true, 4); else DAB_POOL_CASE(false, 4);
#undef DAB_POOL_CASE
}
}  // namespace dab

extern "C" DAB_API int dab_pool_l2norm(const void *hidden, int hidden_dtype, const void *mask, int mask_dtype, float *out,
                               void *out_bf16, float *out_inv, int64_t B, int64_t T, int64_t D, void *stream) {
    if (!out && B > 0) return dab::fail(DAB_E_INVALID, "dab_pool_l2norm: hidden/out must not be NULL");
    return dab::pool_run(hidden, hidden_dtype, mask, mask_dtype, out, out_bf16, out_inv, B, T, D, stream, dab::StoreSink{});
}

extern "C" DAB_API int dab_store_ingest(dab_store *s, const void *hidden, int hidden_dtype, const void *mask, int mask_dtype,
                                        int64_t B, int64_t T, void *stream) {
    using namespace dab;
    if (!s) return fail(DAB_E_INVALID, "dab_store_ingest: store is NULL");
    if (B < 0) return fail(DAB_E_INVALID, "dab_store_ingest: B < 0");
    if (B == 0) return DAB_OK;
    std::lock_guard<std::mutex> g(s->mu);
    if (s->size + B > s->capacity) return fail(DAB_E_CAPACITY, "dab_store_ingest: store is full");
    if (s->size > 0 && s->explicit_ids) return fail(DAB_E_INVALID, "dab_store_ingest: the store holds explicit ids; ingested rows get id_base + row");
    DAB_CUDA_OK(cudaSetDevice(s->device));
    // the kernel's outputs ARE the store: rows [size, size + B), their 1/|row| and ids
    StoreSink sink;
    sink.inv_norm = s->inv_norm + s->size;
    sink.ids = s->ids + s->size;
    sink.id_first = s->id_base + s->size;
    float *out_f32 = nullptr;
    void *out_bf16 = nullptr;
    if (s->dtype == DAB_BF16) out_bf16 = static_cast<__nv_bfloat16 *>(s->rows) + (size_t)s->size * s->dim;
    else out_f32 = static_cast<float *>(s->rows) + (size_t)s->size * s->dim;
    if (int rc = pool_run(hidden, hidden_dtype, mask, mask_dtype, out_f32, out_bf16, nullptr, B, T, s->dim, stream, sink)) return rc;
    s->explicit_ids = false;
    s->size += B;
    return DAB_OK;
}
