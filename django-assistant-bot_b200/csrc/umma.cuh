// umma.cuh - tcgen05 / TMEM / TMA PTX wrappers and tensor-map helpers shared by the tensor-core scan kernels.
#pragma once
#include <cuda.h>
#include <string.h>

#include <mutex>

#include "search.cuh"

namespace dab {
namespace umma {

constexpr int kTileRows = 128;      // rows of the corpus operand one CTA feeds to an MMA (UMMA M per CTA)
constexpr int kBlockK = 64;         // bf16 per 128-byte swizzle row
constexpr int kUmmaK = 16;          // K of one tcgen05.mma for 16-bit inputs
constexpr int kAStageBytes = kTileRows * kBlockK * 2;  // 16 KiB

// ---- PTX wrappers --------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap *map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_out, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_out)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte-swizzled shared-memory operand descriptor (sm_100 UMMA):
//   [0,14) start address >> 4, [16,30) leading byte offset >> 4 (unused for swizzled K-major),
//   [32,46) stride byte offset >> 4 (1024 B between 8-row groups), [46,48) version = 1,
//   [61,64) layout = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(1024u >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// kind::f16 instruction descriptor: D=f32 (bits 4-5 = 1), A=B=bf16 (bits 7-9, 10-12 = 1), both K-major,
// N>>3 at bits 17-22, M>>4 at bits 24-28
__host__ __device__ constexpr uint32_t make_idesc(int n, int m = kTileRows) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
// kind::tf32 descriptor: A=B=tf32 (format code 2); fp32 operands are read with their low 13 mantissa bits dropped
__host__ __device__ constexpr uint32_t make_idesc_tf32(int n, int m = kTileRows) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
// fp32 operands on the tensor cores: K = 8 elements (32 bytes) per instruction
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

// A value every lane holds identically (loaded from shared/global memory), made PROVABLY uniform: the
// result of a warp reduction lives in a uniform register.  Call with the whole warp converged.
__device__ __forceinline__ uint32_t warp_uniform(uint32_t v) { return __reduce_or_sync(0xffffffffu, v); }

// a spin that cannot hang the GPU: a lost arrival traps instead (the host sees a launch failure)
__device__ __forceinline__ void mbar_wait_bounded(uint64_t *bar, uint32_t parity) {
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); spins++) {
        if (spins > (1u << 28)) __trap();
    }
}


__device__ __forceinline__ void append_candidate(const SearchScratch &sc, const uint32_t *__restrict__ allow, int q, uint64_t key) {
    const uint32_t row = key_row(key);
    if (allow != nullptr && !((allow[row >> 5] >> (row & 31)) & 1u)) return;
    if (key > sc.tau_key[q]) {
        const uint32_t pos = atomicAdd(&sc.cand_count[q], 1u);
        if (pos < (uint32_t)kCandCap) sc.cand[(size_t)q * kCandCap + pos] = key;
    }
}

// ---- staging without atomics -------------------------------------------------------------------
// Admitting a candidate used to cost a shared-memory atomic and a 16-way divergent branch ladder per
// hit; at ~1 hit per 1000 scores that was most of the epilogue's time once the MMAs ran at their
// floor.  Here every epilogue warp owns a list (cap 0 = no list, hits go straight to the candidate buffer): the slot of a hit is the warp's running count (a
// warp-uniform register) plus the number of hitting lanes below it (one ballot) - no atomics, no
// divergent ladder; the loop below runs once per "some lane still has a hit", all lanes converged.
struct WarpList {
    uint64_t *key;   // [cap]
    uint32_t *q;     // [cap]
    int cap;
};

__device__ __forceinline__ uint32_t pick16(const uint32_t (&v)[16], int j) {
    uint32_t x = v[0];
#pragma unroll
    for (int i = 1; i < 16; i++) x = j == i ? v[i] : x;
    return x;
}

// Packed fp32 arithmetic of sm_100 (SASS FFMA2 / FMNMX3): two fused multiply-adds per instruction, a 3-input max.
__device__ __forceinline__ void ffma2(float &d0, float &d1, float a0, float a1, float b, float c0, float c1) {
    asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\t"
        "mov.b64 rb, {%4, %4};\n\t"
        "mov.b64 rc, {%5, %6};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
        "mov.b64 {%0, %1}, rd;\n\t}"
        : "=f"(d0), "=f"(d1)
        : "f"(a0), "f"(a1), "f"(b), "f"(c0), "f"(c1));
}
__device__ __forceinline__ float max3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// 16-column slab of an accumulator row (lane = corpus row, column = query).  Returns the new count.
// Nearly every score fails its threshold (~1 in 1000 passes), so the slab is first screened with
// m = max_j (v_j * rinv - tau_j): 8 packed FMAs + 8 three-input maxima + one vote per 512 scores, against a
// multiply, a compare and a mask update for every score.  `ntau_s` holds the NEGATED thresholds.  Only when some
// lane's m comes within 1e-6 of passing does the warp run the exact per-score test below (the screen is
// conservative: the exact test decides).
__device__ __forceinline__ uint32_t epilogue_slab_wl(const uint32_t (&v)[16], int c0, float rinv, long long row,
                                                     const float *tau_s, const float *ntau_s, int q_base,
                                                     const uint32_t *__restrict__ allow, const SearchScratch &sc,
                                                     const WarpList &wl, uint32_t cnt) {
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
        const float4 nt = *reinterpret_cast<const float4 *>(ntau_s + c0 + j);
        float d0, d1, d2, d3;
        ffma2(d0, d1, __uint_as_float(v[j + 0]), __uint_as_float(v[j + 1]), rinv, nt.x, nt.y);
        ffma2(d2, d3, __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]), rinv, nt.z, nt.w);
        m = max3(m, d0, d1);
        m = max3(m, d2, d3);
    }
    // zero rows and rows past the end never rank (rinv = 0, or -0.0 for a tombstone)
    if (!__any_sync(0xffffffffu, m >= -1e-6f && rinv > 0.f)) return cnt;
    uint32_t pass = 0;
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
        const float4 tq = *reinterpret_cast<const float4 *>(tau_s + c0 + j);
        pass |= (__uint_as_float(v[j + 0]) * rinv >= tq.x ? 1u : 0u) << (j + 0);
        pass |= (__uint_as_float(v[j + 1]) * rinv >= tq.y ? 1u : 0u) << (j + 1);
        pass |= (__uint_as_float(v[j + 2]) * rinv >= tq.z ? 1u : 0u) << (j + 2);
        pass |= (__uint_as_float(v[j + 3]) * rinv >= tq.w ? 1u : 0u) << (j + 3);
    }
    if (!(rinv > 0.f)) pass = 0;   // zero rows and rows past the end never rank
    const uint32_t lt = (1u << (threadIdx.x & 31)) - 1u;
    while (__any_sync(0xffffffffu, pass != 0)) {
        const bool has = pass != 0;
        const uint32_t b = __ballot_sync(0xffffffffu, has);
        if (has) {
            const int j = __ffs(pass) - 1;
            pass &= pass - 1;
            const uint64_t key = make_key(__uint_as_float(pick16(v, j)) * rinv, (uint32_t)row);
            const uint32_t slot = cnt + __popc(b & lt);
            if (slot < (uint32_t)wl.cap) {
                wl.key[slot] = key;
                wl.q[slot] = (uint32_t)(q_base + c0 + j);
            } else {
                append_candidate(sc, allow, q_base + c0 + j, key);   // list full (level 0, dense clusters)
            }
        }
        cnt += __popc(b);
    }
    return cnt;
}

// The same slab without a staging list: hits go straight to the candidate buffers, one atomic PER COLUMN of the slab
// (lane j reserves the slots of column j for all hitting lanes at once, 16 reservations in flight together) instead
// of one atomic round trip per hit.  Level 0 admits every row (no threshold yet): 164 k candidates at Q = 64, which
// the per-hit loop turned into 64 dependent atomic round trips per warp and tile - 134 us for a 5 MB launch.
// tau_key_s: exact threshold keys (ties on the score are decided by the row) or nullptr when no threshold exists yet.
__device__ __forceinline__ void epilogue_slab_direct(const uint32_t (&v)[16], int c0, float rinv, long long row,
                                                     const float *tau_s, const float *ntau_s, const uint64_t *tau_key_s,
                                                     int q_base, const uint32_t *__restrict__ allow, const SearchScratch &sc) {
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
        const float4 nt = *reinterpret_cast<const float4 *>(ntau_s + c0 + j);
        float d0, d1, d2, d3;
        ffma2(d0, d1, __uint_as_float(v[j + 0]), __uint_as_float(v[j + 1]), rinv, nt.x, nt.y);
        ffma2(d2, d3, __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]), rinv, nt.z, nt.w);
        m = max3(m, d0, d1);
        m = max3(m, d2, d3);
    }
    if (!__any_sync(0xffffffffu, m >= -1e-6f && rinv > 0.f)) return;
    uint32_t pass = 0;
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
        const float4 tq = *reinterpret_cast<const float4 *>(tau_s + c0 + j);
        pass |= (__uint_as_float(v[j + 0]) * rinv >= tq.x ? 1u : 0u) << (j + 0);
        pass |= (__uint_as_float(v[j + 1]) * rinv >= tq.y ? 1u : 0u) << (j + 1);
        pass |= (__uint_as_float(v[j + 2]) * rinv >= tq.z ? 1u : 0u) << (j + 2);
        pass |= (__uint_as_float(v[j + 3]) * rinv >= tq.w ? 1u : 0u) << (j + 3);
    }
    if (!(rinv > 0.f)) pass = 0;   // zero rows, tombstones and rows past the end never rank
    if (pass != 0 && allow != nullptr && !((allow[row >> 5] >> (row & 31)) & 1u)) pass = 0;
    // how many hits does the busiest lane hold?  Sparse slabs (the usual case once a threshold exists: one hit here
    // and there) take the short per-hit loop; dense ones the per-column reservation.  Measured at Q = 64 over a
    // 12.5M-row shard, per level: per-hit only 136 / 37 / 74 / 319 us, per-column only 20 / 32 / 103 / 413 us.
    const uint32_t busiest = __reduce_max_sync(0xffffffffu, (uint32_t)__popc(pass));
    if (busiest == 0) return;
    if (busiest < 3) {
        while (pass != 0) {
            const int j = __ffs(pass) - 1;
            pass &= pass - 1;
            const int q = q_base + c0 + j;
            const uint64_t key = make_key(__uint_as_float(pick16(v, j)) * rinv, (uint32_t)row);
            if (tau_key_s == nullptr || key > tau_key_s[c0 + j]) {   // ties on the score are decided by the row
                const uint32_t pos = atomicAdd(&sc.cand_count[q], 1u);
                if (pos < (uint32_t)kCandCap) sc.cand[(size_t)q * kCandCap + pos] = key;
            }
        }
        return;
    }
    if (tau_key_s != nullptr) {
#pragma unroll
        for (int j = 0; j < 16; j++)
            if (((pass >> j) & 1u) && !(make_key(__uint_as_float(v[j]) * rinv, (uint32_t)row) > tau_key_s[c0 + j])) pass &= ~(1u << j);
    }
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t col_ballot = 0;       // lane j < 16 keeps which lanes (rows) hit in column j
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const uint32_t b = __ballot_sync(0xffffffffu, (pass >> j) & 1u);
        if (lane == j) col_ballot = b;
    }
    uint32_t base = 0;
    if (lane < 16 && col_ballot != 0) base = atomicAdd(&sc.cand_count[q_base + c0 + lane], (uint32_t)__popc(col_ballot));
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const uint32_t b = __shfl_sync(0xffffffffu, col_ballot, j);
        const uint32_t bj = __shfl_sync(0xffffffffu, base, j);
        if ((pass >> j) & 1u) {
            const uint32_t pos = bj + __popc(b & lt);
            if (pos < (uint32_t)kCandCap)
                sc.cand[(size_t)(q_base + c0 + j) * kCandCap + pos] = make_key(__uint_as_float(v[j]) * rinv, (uint32_t)row);
        }
    }
}

// whole accumulator row over `ncols` columns (multiple of 16, warp-uniform); returns the warp's hit count
// wl.cap == 0: no staging list - hits are written straight to the candidate buffers (epilogue_slab_direct; tau_key_s as there)
__device__ __forceinline__ uint32_t epilogue_tile_wl(uint32_t taddr, int ncols, float rinv, long long row, const float *tau_s,
                                                     const float *ntau_s, int q_base, const uint32_t *__restrict__ allow,
                                                     const SearchScratch &sc, const WarpList &wl, const uint64_t *tau_key_s = nullptr) {
    uint32_t cnt = 0;
    if (ncols <= 0) return cnt;
    uint32_t v0[16], v1[16];
    tmem_ld16(taddr, v0);
    for (int c0 = 0; c0 < ncols; c0 += 32) {
        tmem_ld_wait();
        const bool second = c0 + 16 < ncols;
        if (second) tmem_ld16(taddr + (uint32_t)(c0 + 16), v1);
        if (wl.cap == 0) epilogue_slab_direct(v0, c0, rinv, row, tau_s, ntau_s, tau_key_s, q_base, allow, sc);
        else cnt = epilogue_slab_wl(v0, c0, rinv, row, tau_s, ntau_s, q_base, allow, sc, wl, cnt);
        if (second) {
            tmem_ld_wait();
            if (c0 + 32 < ncols) tmem_ld16(taddr + (uint32_t)(c0 + 32), v0);
            if (wl.cap == 0) epilogue_slab_direct(v1, c0 + 16, rinv, row, tau_s, ntau_s, tau_key_s, q_base, allow, sc);
            else cnt = epilogue_slab_wl(v1, c0 + 16, rinv, row, tau_s, ntau_s, q_base, allow, sc, wl, cnt);
        }
    }
    return cnt;
}

// The flusher warp's side: `nlists` (<= 8) lists of `cap` entries laid out back to back, counts in cnt_s.
// One flattened pass, so a tile's candidates cost one global round trip, not one per list.
__device__ __forceinline__ void flush_warp_lists(const uint64_t *key, const uint32_t *q, const uint32_t *cnt_s, int nlists, int cap,
                                                 const uint32_t *__restrict__ allow, const SearchScratch &sc) {
    const int lane = threadIdx.x & 31;
    uint32_t incl = lane < nlists ? min(cnt_s[lane], (uint32_t)cap) : 0u;
#pragma unroll
    for (int d = 1; d < 8; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    uint32_t end[8];
#pragma unroll
    for (int w = 0; w < 8; w++) end[w] = __shfl_sync(0xffffffffu, incl, w);
    const uint32_t total = end[7];
    for (uint32_t i = lane; i < total; i += 32) {
        int w = 0;
        uint32_t begin = 0;
#pragma unroll
        for (int ww = 0; ww < 7; ww++)
            if (i >= end[ww]) {
                w = ww + 1;
                begin = end[ww];
            }
        const uint32_t e = (uint32_t)(w * cap) + (i - begin);
        append_candidate(sc, allow, (int)q[e], key[e]);
    }
}

// ---- 2-CTA (cta_group::2) and cluster helpers ------------------------------------------------
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // clears the pair bit of a shared::cluster address -> even CTA of the pair

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t *bar, uint32_t rank) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
        ::"r"(smem_u32(bar)), "r"(rank)
        : "memory");
}
// A-tile load, multicast to the CTAs in `mask`; complete_tx goes to the pair LEADER of every destination
__device__ __forceinline__ void tma_load_2d_mc2(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%4, %5}], [%2], %3;"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "h"(mask), "r"(c0), "r"(c1)
        : "memory");
}
// A-tile load into THIS CTA only; complete_tx goes to the pair leader's mbarrier
__device__ __forceinline__ void tma_load_2d_2cta(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t *smem_out, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_out)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_f16_2cta(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_tf32_2cta(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive (once) on the mbarrier at this offset in every CTA of `mask` when the MMAs issued so far are done
__device__ __forceinline__ void umma_commit_mc2(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask)
                 : "memory");
}


__device__ __forceinline__ long long level_tile(const LevelPlan &plan, int lvl, long long j) {
    const long long i = lvl == 0 ? j : j + j / (plan.growth - 1) + 1;
    return i * plan.stride[lvl];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 2-D row-major [rows, dim] tensor (bf16, or fp32 when f32 is set), box = [box_rows x 128 bytes of columns],
// 128-byte swizzle, OOB rows read as zero
inline int make_map(CUtensorMap *map, const void *base, uint64_t rows, uint64_t dim, uint32_t box_rows, bool f32 = false) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return fail(DAB_E_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    const cuuint64_t gdim[2] = {dim, rows};
    const cuuint64_t gstride[1] = {dim * (f32 ? 4u : 2u)};
    const cuuint32_t box[2] = {(cuuint32_t)(f32 ? kBlockK / 2 : kBlockK), box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                          const_cast<void *>(base), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(DAB_E_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return DAB_OK;
}

// ---- host work kept off the per-launch path -----------------------------------------------------
// A search enqueues one scan launch per level; encoding two tensor maps, setting the shared-memory
// attribute and asking for the cluster occupancy on every one of them was measurable on small stores
// (1M x 768 fp32, Q=1: 40 us of a 490 us search).  All three are pure functions of their arguments.

// Tensor maps by (base, rows, dim, box, type): a small round-robin cache shared by the scan launchers.
inline int cached_map(CUtensorMap *out, const void *base, uint64_t rows, uint64_t dim, uint32_t box_rows, bool f32 = false) {
    struct Entry {
        const void *base;
        uint64_t rows, dim;
        uint32_t box_rows;
        bool f32, valid;
        CUtensorMap map;
    };
    constexpr int kEntries = 64;
    static Entry cache[kEntries] = {};
    static int next = 0;
    static std::mutex mu;
    {
        std::lock_guard<std::mutex> g(mu);
        for (int i = 0; i < kEntries; i++) {
            const Entry &e = cache[i];
            if (e.valid && e.base == base && e.rows == rows && e.dim == dim && e.box_rows == box_rows && e.f32 == f32) {
                memcpy(out, &e.map, sizeof(CUtensorMap));
                return DAB_OK;
            }
        }
    }
    {
        // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
        // GPU's memory hung the GPU: a tensor map is only ever built over memory of the current device.
        cudaPointerAttributes pa;
        int dev = -1;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaPointerGetAttributes(&pa, base) != cudaSuccess ||
            pa.type != cudaMemoryTypeDevice || pa.device != dev)
            return fail(DAB_E_INVALID, "tensor map requested over memory that is not device memory of the current GPU");
    }
    if (int rc = make_map(out, base, rows, dim, box_rows, f32)) return rc;
    std::lock_guard<std::mutex> g(mu);
    Entry &e = cache[next];
    next = (next + 1) % kEntries;
    e.base = base; e.rows = rows; e.dim = dim; e.box_rows = box_rows; e.f32 = f32; e.valid = true;
    memcpy(&e.map, out, sizeof(CUtensorMap));
    return DAB_OK;
}

// cudaOccupancyMaxActiveClusters by (device, shared memory, cluster size), one instance per kernel.
struct ClusterOccupancyCache {
    struct Entry { int dev; size_t smem; int cluster; int value; };
    Entry e[16] = {};
    int n = 0;
    std::mutex mu;
    template <typename K>
    cudaError_t get(int *out, K kernel, cudaLaunchConfig_t *cfg, int cluster) {
        int dev = 0;
        cudaError_t err = cudaGetDevice(&dev);
        if (err != cudaSuccess) return err;
        {
            std::lock_guard<std::mutex> g(mu);
            for (int i = 0; i < n; i++)
                if (e[i].dev == dev && e[i].smem == cfg->dynamicSmemBytes && e[i].cluster == cluster) {
                    *out = e[i].value;
                    return cudaSuccess;
                }
        }
        err = cudaOccupancyMaxActiveClusters(out, kernel, cfg);
        if (err != cudaSuccess) return err;
        std::lock_guard<std::mutex> g(mu);
        if (n < 16) e[n++] = Entry{dev, cfg->dynamicSmemBytes, cluster, *out};
        return cudaSuccess;
    }
};

}  // namespace umma
}  // namespace dab
