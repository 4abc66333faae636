// Batched residual correlation on the 5th-generation tensor cores (sm_100a):
//     Q[r][n] = sum_m At[n][m] * P[r][m]        (the reference's x gemv(A^T, .) calls, homotopy-cpu.cpp:97,:120,
//                                                batched over signals: one dense A^T . P contraction)
// float32 accuracy comes from the 3xTF32 split  a*p ~= a_hi*p_hi + (a_lo*p_hi + a_hi*p_lo)  with
// hi = rna_tf32(v), lo = rna_tf32(v - hi): three tcgen05.mma.kind::tf32 per K-step.  The tensor core adds
// into its fp32 accumulator with round-toward-zero, a bias that grows with the number of accumulation
// events (measured: -2.5e-6 relative with all three products in one accumulator at M = 512, gpurun_out/
// r1_contract_c2.json).  The two small correction products therefore get their OWN accumulator and are
// added once, round-to-nearest, in the epilogue: the main accumulator sees one event per K-step instead of 3.
//
// Kernel shape (persistent, one CTA per SM, 192 threads):
//   warp 0      TMA producer: per K-block (32 floats = one 128-byte swizzle span) loads four tiles
//               A_hi, A_lo [128 atoms x 32] and P_hi, P_lo [256 signals x 32] (cp.async.bulk.tensor.2d, SWIZZLE_128B)
//   warp 1      MMA issuer (one lane): UMMA M=128 (atoms) x N=256 (signals) x K=8; TMEM columns [0,256) hold
//               the main accumulator, [256,512) the correction accumulator (all 512 columns of the SM)
//   warps 2..9  epilogue: tcgen05.ld 32 lanes x 32 columns of both accumulators -> registers -> add -> Q;
//               lane = atom, so every store instruction writes 128 contiguous bytes of one signal's row
// Barriers: full/empty per smem stage (TMA <-> MMA), tmem_full/tmem_empty for the accumulator pair (MMA <-> epilogue).
#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "solver_base.h"

namespace ssb {

namespace {

constexpr int TILE_ATOMS = 128;     // UMMA M
constexpr int TILE_SIGS = 256;      // UMMA N
// K-block = one swizzle span of every operand row.  128-byte rows (64 fp16 / 32 tf32 elements): three 64 KB stages per
// CTA in the pair kernel; 64-byte rows (-DSSB_BLOCK_K=16): six 32 KB stages.  Measured on C2 the two are within 3 %
// (gpurun_out bench A/B, 368/359 vs 353/354 TFLOP/s), the wider rows ahead: fewer, larger TMA boxes and barriers.
#ifndef SSB_BLOCK_K
#define SSB_BLOCK_K 32
#endif
constexpr int BLOCK_K = SSB_BLOCK_K;   // floats per K-block: 32 = one 128-byte swizzle span, 16 = one 64-byte span
constexpr int ROW_BYTES = BLOCK_K * 4;
constexpr int UMMA_K = 8;           // tf32
constexpr int STAGES = BLOCK_K == 32 ? 2 : 4;
constexpr int A_TILE_BYTES = TILE_ATOMS * BLOCK_K * 4;   // 16 KB
constexpr int P_TILE_BYTES = TILE_SIGS * BLOCK_K * 4;    // 32 KB
constexpr int STAGE_BYTES = 2 * A_TILE_BYTES + 2 * P_TILE_BYTES;   // 96 KB
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;  // + alignment slack
constexpr int TMEM_COLS = 512;
constexpr int EPI_WARPS = 8;
constexpr int THREADS = 64 + 32 * EPI_WARPS;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d_mc(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
        : "memory");
}
// arrives on the barrier at the same offset in every CTA of `mask` once the MMAs issued so far have retired
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_size() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }

// ---- CTA-pair (cta_group::2) variants: one UMMA spans two SMs, M = 256 atoms, each SM feeds its own 128 atom rows and
// HALF of the signal rows from its own shared memory -- the operand bytes each SM moves per MAC drop by a third.
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint64_t* leader_bar, int c0, int c1) {
    // executed by both CTAs; clearing the peer bit makes the transaction bytes land on CTA 0's barrier
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(leader_bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {   // arrives on `bar` in both CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cta0(uint64_t* bar) {    // arrive on CTA 0's copy of `bar`
    asm volatile(
        "{\n"
        ".reg .b32 ra;\n"
        "mapa.shared::cluster.u32 ra, %0, 0;\n"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
        "}\n" ::"r"(smem_u32(bar)) : "memory");
}
// one wrapper per (container type, CTA group): F16 = operands are fp16 (kind::f16), else tf32 containers (kind::tf32)
template <bool F16, bool PAIR>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if constexpr (F16 && PAIR)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                     ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    else if constexpr (F16)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                     ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    else if constexpr (PAIR) umma_tf32_pair(tmem_d, adesc, bdesc, idesc, accumulate);
    else umma_tf32(tmem_d, adesc, bdesc, idesc, accumulate);
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}

// K-major operand tile, rows of 128 bytes, SWIZZLE_128B: 8-row groups are 1024 bytes apart
__device__ __forceinline__ uint64_t make_desc(const void* tile) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_u32(tile) >> 4) & 0x3FFF);   // start address
    d |= (uint64_t)((8 * ROW_BYTES) >> 4) << 32;       // stride byte offset: one 8-row swizzle group
    d |= (uint64_t)1 << 46;                            // descriptor version (sm_100)
    d |= (uint64_t)(ROW_BYTES == 128 ? 2 : 4) << 61;   // SWIZZLE_128B / SWIZZLE_64B
    return d;
}

// instruction descriptor: D = F32, A = B = TF32, both K-major, M = 128, N = 256
// (a/b format: 0 = F16 under kind::f16, 2 = TF32 under kind::tf32)
__host__ __device__ constexpr uint32_t make_idesc(int n_sigs, bool f16, int m_atoms = TILE_ATOMS) {
    return (1u << 4) | ((f16 ? 0u : 2u) << 7) | ((f16 ? 0u : 2u) << 10) | ((uint32_t)(n_sigs >> 3) << 17) | ((uint32_t)(m_atoms >> 4) << 24);
}
__device__ int g_tc_dbg = 0;   // timing experiments only (SSB200_TC_DBG): 1 no Q stores, 2 no MMA issue, 4 no TMEM drain
constexpr float CORR_SCALE = 1.0f / 2048.0f;   // the fp16 lo parts are stored times 2^11
constexpr int THIN_SIGS = 32;       // UMMA N of the thin-tail variant (same accumulation per element, 1/8 of the tensor and TMA work)

template <bool F16>
__global__ void __launch_bounds__(THREADS, 1)
contract_x3_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                   const __grid_constant__ CUtensorMap map_p_hi, const __grid_constant__ CUtensorMap map_p_lo,
                   float* __restrict__ Q, int ldq, int rows, int atom_tiles, int sig_tiles, int k_blocks, int tile_sigs,
                   const float* __restrict__ a_inv, const float* __restrict__ p_inv, const int* __restrict__ rows_dev,
                   int nblk, int blk_off)
{
    constexpr int KB_ELEMS = F16 ? ROW_BYTES / 2 : ROW_BYTES / 4;     // operand elements per K-block
    if (rows_dev) { rows = min(rows, *rows_dev); sig_tiles = (rows + tile_sigs - 1) / tile_sigs; }   // host count may be stale (larger)
    // nblk row blocks of `rows` rows each, blk_off rows apart in P and Q (exact-residual mode: block 0 = P, block 1 = R)
    const int blk_tiles = sig_tiles;
    sig_tiles *= nblk;
    auto tile_row0 = [&](int st) { return (st / blk_tiles) * blk_off + (st % blk_tiles) * tile_sigs; };
    auto tile_end = [&](int st) { return (st / blk_tiles) * blk_off + rows; };
    extern __shared__ uint8_t smem_raw[];
    constexpr int MAX_STAGES = 8;
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES], empty_bar[MAX_STAGES], tmem_full_bar, tmem_empty_bar;
    __shared__ uint32_t tmem_base_slot;

    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // stage = A_hi | A_lo | P_hi | P_lo for tile_sigs signal rows: the thin tile leaves room for more stages, and the
    // tail iterations it serves are pure latency (16 K-blocks, one CTA per atom tile)
    const int p_bytes = tile_sigs * ROW_BYTES;
    const int stage_bytes = 2 * A_TILE_BYTES + 2 * p_bytes;
    const int nstages = min(MAX_STAGES, (STAGES * STAGE_BYTES) / stage_bytes);
    // A cluster of `cs` CTAs works on cs neighbouring atom tiles of ONE signal tile: each CTA fetches 1/cs of the
    // P tile and multicasts it to all of them, so the signal operand crosses the L2 -> SM fabric once per cluster
    // (the operand stream, not the tensor pipe, is what bounded the single-CTA kernel: 61 of ~64 B/clk/SM).
    const int cs = (int)cluster_size(), crank = (int)cluster_rank();
    const uint16_t cmask = (uint16_t)((1u << cs) - 1u);
    const int num_items = (atom_tiles / cs) * sig_tiles;          // work items of one cluster-wide step
    const int cluster_id = blockIdx.x / cs, num_clusters = gridDim.x / cs;

    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], cs); }
        mbar_init(&tmem_full_bar, 1);
        mbar_init(&tmem_empty_bar, EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (cs > 1) cluster_sync_all();       // every CTA's barriers exist before a peer multicasts into them
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            uint32_t it = 0;
            const int part_rows = tile_sigs / cs;                       // rows of the P tile this CTA fetches
            const int part_bytes = part_rows * ROW_BYTES;
            for (int item = cluster_id; item < num_items; item += num_clusters) {
                const int n0 = ((item % (atom_tiles / cs)) * cs + crank) * TILE_ATOMS;
                const int r0 = tile_row0(item / (atom_tiles / cs));
                for (int kb = 0; kb < k_blocks; ++kb, ++it) {
                    const int s = it % nstages;
                    mbar_wait(&empty_bar[s], ((it / nstages) & 1) ^ 1);      // every CTA of the cluster has drained stage s
                    uint8_t* st = smem + (size_t)s * stage_bytes;
                    mbar_expect_tx(&full_bar[s], 2 * A_TILE_BYTES + 2 * tile_sigs * ROW_BYTES);
                    tma_load_2d(st, &map_a_hi, &full_bar[s], kb * KB_ELEMS, n0);
                    tma_load_2d(st + A_TILE_BYTES, &map_a_lo, &full_bar[s], kb * KB_ELEMS, n0);
                    uint8_t* ph = st + 2 * A_TILE_BYTES + crank * part_bytes;
                    if (cs == 1) {
                        tma_load_2d(ph, &map_p_hi, &full_bar[s], kb * KB_ELEMS, r0);
                        tma_load_2d(ph + p_bytes, &map_p_lo, &full_bar[s], kb * KB_ELEMS, r0);
                    } else {
                        tma_load_2d_mc(ph, &map_p_hi, &full_bar[s], kb * KB_ELEMS, r0 + crank * part_rows, cmask);
                        tma_load_2d_mc(ph + p_bytes, &map_p_lo, &full_bar[s], kb * KB_ELEMS, r0 + crank * part_rows, cmask);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            uint32_t it = 0, t = 0;
            const uint32_t d_main = tmem_base, d_corr = tmem_base + TILE_SIGS;
            const uint32_t IDESC = make_idesc(tile_sigs, F16);
            for (int item = cluster_id; item < num_items; item += num_clusters, ++t) {
                mbar_wait(&tmem_empty_bar, (t & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int kb = 0; kb < k_blocks; ++kb, ++it) {
                    const int s = it % nstages;
                    mbar_wait(&full_bar[s], (it / nstages) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    uint8_t* st = smem + (size_t)s * stage_bytes;
                    const uint64_t a_hi = make_desc(st), a_lo = make_desc(st + A_TILE_BYTES);
                    const uint64_t p_hi = make_desc(st + 2 * A_TILE_BYTES), p_lo = make_desc(st + 2 * A_TILE_BYTES + p_bytes);
#pragma unroll
                    for (int kk = 0; kk < ROW_BYTES / 32; ++kk) {                  // one UMMA K-step = 32 bytes of every row
                        const uint64_t adv = (uint64_t)(kk * 2);                   // ... in 16-byte units inside the swizzle span
                        umma<F16, false>(d_corr, a_lo + adv, p_hi + adv, IDESC, (kb | kk) != 0);
                        umma<F16, false>(d_corr, a_hi + adv, p_lo + adv, IDESC, 1);
                        umma<F16, false>(d_main, a_hi + adv, p_hi + adv, IDESC, (kb | kk) != 0);
                    }
                    if (cs == 1) umma_commit(&empty_bar[s]);   // frees the smem stage when these MMAs retire
                    else umma_commit_mc(&empty_bar[s], cmask);  // ... in every CTA that multicasts into it
                }
                umma_commit(&tmem_full_bar);               // both accumulators complete
            }
        }
    } else {
        // ===== epilogue: TMEM -> registers -> Q =====
        const int quarter = warp & 3;                      // TMEM lanes this warp may read
        const int half = (warp - 2) >> 2;                  // which 128 signal columns of the tile
        uint32_t t = 0;
        for (int item = cluster_id; item < num_items; item += num_clusters, ++t) {
            const int n0 = ((item % (atom_tiles / cs)) * cs + crank) * TILE_ATOMS;
            const int r0 = tile_row0(item / (atom_tiles / cs)) + half * (TILE_SIGS / 2);
            const int r_end = tile_end(item / (atom_tiles / cs));
            mbar_wait(&tmem_full_bar, t & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * (TILE_SIGS / 2));
            float* qcol = Q + n0 + quarter * 32 + lane;
            const float ainv = F16 ? __ldg(a_inv + n0 + quarter * 32 + lane) : 1.0f;   // undoes the atom's power-of-two scale
            // drain both accumulators into registers first and hand TMEM back: the next tile's MMAs run while
            // this tile's stores are still in flight (with both accumulators resident TMEM cannot be double-buffered)
            float r[TILE_SIGS / 2];
#pragma unroll
            for (int c0 = 0; c0 < TILE_SIGS / 2; c0 += 32) {
                uint32_t v[32], w[32];
                tmem_ld32(taddr0 + (uint32_t)c0, v);
                tmem_ld32(taddr0 + (uint32_t)(TILE_SIGS + c0), w);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    r[c0 + j] = F16 ? fmaf(__uint_as_float(w[j]), CORR_SCALE, __uint_as_float(v[j])) : __uint_as_float(v[j]) + __uint_as_float(w[j]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar);
#pragma unroll
            for (int j = 0; j < TILE_SIGS / 2; ++j) {
                const int rr = r0 + j;
                if (rr < r_end && half * (TILE_SIGS / 2) + j < tile_sigs) qcol[(size_t)rr * ldq] = F16 ? (r[j] * ainv) * __ldg(p_inv + rr) : r[j];
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (cs > 1) cluster_sync_all();       // no CTA leaves while a peer can still write its shared memory or barriers
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

// ---- CTA-pair kernel: the tile is 256 atoms x 256 signals over two SMs (cluster of 2, same TPC).  CTA r owns atom rows
// [128 r, 128 r + 128) of the tile (operand A and the matching 128 TMEM lanes) and signal rows [128 r, 128 r + 128) of
// the P tile (its half of operand B).  Both CTAs run a TMA producer whose bytes land on CTA 0's full barrier; CTA 0
// alone issues tcgen05.mma.cta_group::2; its commits arrive on both CTAs' empty / tmem_full barriers; both CTAs'
// epilogue warps drain their own TMEM half and arrive on CTA 0's tmem_empty barrier.
constexpr int PAIR_HALF_SIGS = TILE_SIGS / 2;
constexpr int PAIR_P_BYTES = PAIR_HALF_SIGS * ROW_BYTES;                  // 8 KB
constexpr int PAIR_STAGE_BYTES = 2 * A_TILE_BYTES + 2 * PAIR_P_BYTES;     // 32 KB per CTA
constexpr int PAIR_STAGES = (192 * 1024) / PAIR_STAGE_BYTES;   // 6 with 64-byte rows
constexpr int PAIR_SMEM_BYTES = PAIR_STAGES * PAIR_STAGE_BYTES + 1024;

template <bool F16>
__global__ void __launch_bounds__(THREADS, 1)
contract_x3_pair_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                        const __grid_constant__ CUtensorMap map_p_hi, const __grid_constant__ CUtensorMap map_p_lo,
                        float* __restrict__ Q, int ldq, int rows, int atom_pairs, int sig_tiles, int k_blocks,
                        const float* __restrict__ a_inv, const float* __restrict__ p_inv, const int* __restrict__ rows_dev,
                        int nblk, int blk_off)
{
    constexpr int KB_ELEMS = F16 ? ROW_BYTES / 2 : ROW_BYTES / 4;
    if (rows_dev) { rows = min(rows, *rows_dev); sig_tiles = (rows + TILE_SIGS - 1) / TILE_SIGS; }
    // nblk row blocks of `rows` rows each, blk_off rows apart in P and Q (exact-residual mode: block 0 = P, block 1 = R)
    const int blk_tiles = sig_tiles;
    sig_tiles *= nblk;
    auto tile_row0 = [&](int st) { return (st / blk_tiles) * blk_off + (st % blk_tiles) * TILE_SIGS; };
    auto tile_end = [&](int st) { return (st / blk_tiles) * blk_off + rows; };
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[PAIR_STAGES], empty_bar[PAIR_STAGES], tmem_full_bar, tmem_empty_bar;
    __shared__ uint32_t tmem_base_slot;

    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int crank = (int)cluster_rank();
    const int dbg = g_tc_dbg;
    const int num_items = atom_pairs * sig_tiles;
    const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < PAIR_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&tmem_full_bar, 1);
        mbar_init(&tmem_empty_bar, 2 * EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (lane == 0) {
            uint32_t it = 0;
            for (int item = pair_id; item < num_items; item += num_pairs) {
                const int n0 = ((item % atom_pairs) * 2 + crank) * TILE_ATOMS;
                const int r0 = tile_row0(item / atom_pairs) + crank * PAIR_HALF_SIGS;
                for (int kb = 0; kb < k_blocks; ++kb, ++it) {
                    const int s = it % PAIR_STAGES;
                    mbar_wait(&empty_bar[s], ((it / PAIR_STAGES) & 1) ^ 1);
                    uint8_t* st = smem + (size_t)s * PAIR_STAGE_BYTES;
                    if (crank == 0) mbar_expect_tx(&full_bar[s], 2 * PAIR_STAGE_BYTES);   // both CTAs' bytes
                    tma_load_2d_pair(st, &map_a_hi, &full_bar[s], kb * KB_ELEMS, n0);
                    tma_load_2d_pair(st + A_TILE_BYTES, &map_a_lo, &full_bar[s], kb * KB_ELEMS, n0);
                    tma_load_2d_pair(st + 2 * A_TILE_BYTES, &map_p_hi, &full_bar[s], kb * KB_ELEMS, r0);
                    tma_load_2d_pair(st + 2 * A_TILE_BYTES + PAIR_P_BYTES, &map_p_lo, &full_bar[s], kb * KB_ELEMS, r0);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (CTA 0 only) =====
        if (lane == 0 && crank == 0) {
            uint32_t it = 0, t = 0;
            const uint32_t d_main = tmem_base, d_corr = tmem_base + TILE_SIGS;
            const uint32_t IDESC = make_idesc(TILE_SIGS, F16, 2 * TILE_ATOMS);   // M = 256 over the pair
            for (int item = pair_id; item < num_items; item += num_pairs, ++t) {
                mbar_wait(&tmem_empty_bar, (t & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int kb = 0; kb < k_blocks; ++kb, ++it) {
                    const int s = it % PAIR_STAGES;
                    mbar_wait(&full_bar[s], (it / PAIR_STAGES) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    uint8_t* st = smem + (size_t)s * PAIR_STAGE_BYTES;
                    const uint64_t a_hi = make_desc(st), a_lo = make_desc(st + A_TILE_BYTES);
                    const uint64_t p_hi = make_desc(st + 2 * A_TILE_BYTES), p_lo = make_desc(st + 2 * A_TILE_BYTES + PAIR_P_BYTES);
#pragma unroll
                    for (int kk = 0; kk < ROW_BYTES / 32; ++kk) {
                        const uint64_t adv = (uint64_t)(kk * 2);
                        if (dbg & 2) continue;
                        umma<F16, true>(d_corr, a_lo + adv, p_hi + adv, IDESC, (kb | kk) != 0);
                        umma<F16, true>(d_corr, a_hi + adv, p_lo + adv, IDESC, 1);
                        umma<F16, true>(d_main, a_hi + adv, p_hi + adv, IDESC, (kb | kk) != 0);
                    }
                    umma_commit_pair(&empty_bar[s]);
                }
                umma_commit_pair(&tmem_full_bar);
            }
        }
    } else {
        // ===== epilogue (both CTAs): own TMEM half -> registers -> Q =====
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;
        uint32_t t = 0;
        for (int item = pair_id; item < num_items; item += num_pairs, ++t) {
            const int n0 = ((item % atom_pairs) * 2 + crank) * TILE_ATOMS;
            const int r0 = tile_row0(item / atom_pairs) + half * (TILE_SIGS / 2);
            const int r_end = tile_end(item / atom_pairs);
            mbar_wait(&tmem_full_bar, t & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * (TILE_SIGS / 2));
            float* qcol = Q + n0 + quarter * 32 + lane;
            const float ainv = F16 ? __ldg(a_inv + n0 + quarter * 32 + lane) : 1.0f;   // undoes the atom's power-of-two scale
            float r[TILE_SIGS / 2];
#pragma unroll
            for (int c0 = 0; c0 < TILE_SIGS / 2; c0 += 32) {
                if (dbg & 4) { for (int j = 0; j < 32; ++j) r[c0 + j] = 0.f; continue; }
                uint32_t v[32], w[32];
                tmem_ld32(taddr0 + (uint32_t)c0, v);
                tmem_ld32(taddr0 + (uint32_t)(TILE_SIGS + c0), w);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    r[c0 + j] = F16 ? fmaf(__uint_as_float(w[j]), CORR_SCALE, __uint_as_float(v[j])) : __uint_as_float(v[j]) + __uint_as_float(w[j]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive_cta0(&tmem_empty_bar);
#pragma unroll
            for (int j = 0; j < TILE_SIGS / 2; ++j) {
                const int rr = r0 + j;
                if (rr < r_end && !(dbg & 1)) qcol[(size_t)rr * ldq] = F16 ? (r[j] * ainv) * __ldg(p_inv + rr) : r[j];
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

// ---- TMA descriptors through the driver entry point (no link-time libcuda dependency) -------------------
using encode_fn_t = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_fn_t get_encode()
{
    static encode_fn_t fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<encode_fn_t>(p);
    });
    return fn;
}

CUtensorMap make_map(bool f16, const void* base, int inner, int outer, int box_rows)
{
    encode_fn_t enc = get_encode();
    if (!enc) throw invalid_argument{ "cuTensorMapEncodeTiled is not available from this driver" };
    CUtensorMap m;
    const size_t esz = f16 ? 2 : 4;
    cuuint64_t dims[2] = { (cuuint64_t)inner, (cuuint64_t)outer };
    cuuint64_t strides[1] = { (cuuint64_t)inner * esz };
    cuuint32_t box[2] = { (cuuint32_t)(ROW_BYTES / esz), (cuuint32_t)box_rows };
    cuuint32_t estr[2] = { 1, 1 };
    CUresult r = enc(&m, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims,
                     strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     ROW_BYTES == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) throw invalid_argument{ "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")" };
    return m;
}

template <bool F16>
void launch(const TcPlan* p, int rows, float* Q, int ldq, cudaStream_t stream, const int* rows_dev, int nblk, int blk_off);

}  // namespace

struct TcPlan {
    bool f16;                                 // operand containers: fp16 with power-of-two row scales, or tf32
    CUtensorMap a_hi, a_lo, p_hi, p_lo, p_hi_thin, p_lo_thin;
    CUtensorMap p_hi_part[3], p_lo_part[3];   // 1/2 and 1/4 of the signal tile per box (index = log2 cluster size)
    const float* a_inv; const float* p_inv;   // fp16 containers: 1 / scale of every atom row and every signal row
    int cluster;                              // single-CTA kernel: CTAs that share one multicast P tile
    bool pair;                                // CTA-pair (cta_group::2) kernel for full-width tiles
    int n_pad, mp, rows_alloc, sms;
};

bool tc_supported(int n_pad, int mp) { return n_pad % TILE_ATOMS == 0 && mp % 32 == 0 && get_encode() != nullptr; }

TcPlan* tc_plan_create(bool f16, const void* At_hi, const void* At_lo, int n_pad, int mp, const void* P_hi, const void* P_lo,
                       int rows_alloc, const float* a_inv, const float* p_inv)
{
    SSB_REQUIRE(rows_alloc % TILE_SIGS == 0, "P buffers must be allocated in multiples of 256 rows");
    auto* p = new TcPlan;
    try {
        p->a_hi = make_map(f16, At_hi, mp, n_pad, TILE_ATOMS);
        p->a_lo = make_map(f16, At_lo, mp, n_pad, TILE_ATOMS);
        p->p_hi = make_map(f16, P_hi, mp, rows_alloc, TILE_SIGS);
        p->p_lo = make_map(f16, P_lo, mp, rows_alloc, TILE_SIGS);
        p->p_hi_thin = make_map(f16, P_hi, mp, rows_alloc, THIN_SIGS);
        p->p_lo_thin = make_map(f16, P_lo, mp, rows_alloc, THIN_SIGS);
        for (int l = 1; l <= 2; ++l) {
            p->p_hi_part[l] = make_map(f16, P_hi, mp, rows_alloc, TILE_SIGS >> l);
            p->p_lo_part[l] = make_map(f16, P_lo, mp, rows_alloc, TILE_SIGS >> l);
        }
    } catch (...) { delete p; throw; }
    p->f16 = f16; p->a_inv = a_inv; p->p_inv = p_inv;
    { int dbg = 0; if (const char* e = getenv("SSB200_TC_DBG")) dbg = atoi(e); SSB_CUDA(cudaMemcpyToSymbol(g_tc_dbg, &dbg, sizeof(int))); }
    p->n_pad = n_pad; p->mp = mp; p->rows_alloc = rows_alloc;
    p->cluster = 1;
    if (const char* e = getenv("SSB200_CLUSTER")) p->cluster = atoi(e);                              // tuning knob: 1, 2 or 4
    if (p->cluster != 1 && p->cluster != 2 && p->cluster != 4) p->cluster = 1;
    while (p->cluster > 1 && (n_pad / TILE_ATOMS) % p->cluster != 0) p->cluster >>= 1;
    p->pair = (n_pad / TILE_ATOMS) % 2 == 0;
    if (const char* e = getenv("SSB200_PAIR")) p->pair = p->pair && atoi(e) != 0;                   // tuning knob
    int dev = 0;
    SSB_CUDA(cudaGetDevice(&dev));
    SSB_CUDA(cudaDeviceGetAttribute(&p->sms, cudaDevAttrMultiProcessorCount, dev));
    SSB_CUDA(cudaFuncSetAttribute(contract_x3_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_BYTES));
    SSB_CUDA(cudaFuncSetAttribute(contract_x3_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_BYTES));
    SSB_CUDA(cudaFuncSetAttribute(contract_x3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    SSB_CUDA(cudaFuncSetAttribute(contract_x3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    return p;
}

void tc_plan_destroy(TcPlan* p) { delete p; }

namespace {
template <bool F16>
void launch(const TcPlan* p, int rows, float* Q, int ldq, cudaStream_t stream, const int* rows_dev, int nblk, int blk_off)
{
    const int atom_tiles = p->n_pad / TILE_ATOMS;
    const bool thin = rows <= THIN_SIGS;
    const int tile_sigs = thin ? THIN_SIGS : TILE_SIGS;
    const int sig_tiles = (rows + tile_sigs - 1) / tile_sigs;
    const int kb_elems = F16 ? ROW_BYTES / 2 : ROW_BYTES / 4;
    const int k_blocks = (p->mp + kb_elems - 1) / kb_elems;   // the last box may reach past Mp: TMA fills it with zeros
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    cfg.blockDim = dim3(THREADS); cfg.stream = stream;
    if (!thin && p->pair) {
        const int atom_pairs = atom_tiles / 2, items = atom_pairs * sig_tiles * nblk;
        cfg.gridDim = dim3((unsigned)(std::min(items, p->sms / 2) * 2));
        cfg.dynamicSmemBytes = PAIR_SMEM_BYTES;
        attr[0].val.clusterDim.x = 2;
        SSB_CUDA(cudaLaunchKernelEx(&cfg, contract_x3_pair_kernel<F16>, p->a_hi, p->a_lo, p->p_hi_part[1], p->p_lo_part[1], Q, ldq,
                                    rows, atom_pairs, sig_tiles, k_blocks, p->a_inv, p->p_inv, rows_dev, nblk, blk_off));
        return;
    }
    const int cs = thin ? 1 : p->cluster, lg = cs == 4 ? 2 : cs == 2 ? 1 : 0;
    const int items = (atom_tiles / cs) * sig_tiles * nblk;
    const CUtensorMap& mh = thin ? p->p_hi_thin : lg ? p->p_hi_part[lg] : p->p_hi;
    const CUtensorMap& ml = thin ? p->p_lo_thin : lg ? p->p_lo_part[lg] : p->p_lo;
    cfg.gridDim = dim3((unsigned)(std::min(items, p->sms / cs) * cs));
    cfg.dynamicSmemBytes = SMEM_BYTES;
    attr[0].val.clusterDim.x = (unsigned)cs;
    SSB_CUDA(cudaLaunchKernelEx(&cfg, contract_x3_kernel<F16>, p->a_hi, p->a_lo, mh, ml, Q, ldq, rows, atom_tiles, sig_tiles,
                                k_blocks, tile_sigs, p->a_inv, p->p_inv, rows_dev, nblk, blk_off));
}
}  // namespace

void tc_contract(TcPlan* p, int rows, float* Q, int ldq, cudaStream_t stream, const int* rows_dev, int nblk, int blk_off)
{
    if (rows <= 0) return;
    SSB_REQUIRE(nblk >= 1 && (nblk - 1) * blk_off + rows <= p->rows_alloc, "more rows than the plan was built for");
    SSB_REQUIRE(nblk == 1 || blk_off % TILE_SIGS == 0, "row blocks must start on a tile boundary");
    if (p->f16) launch<true>(p, rows, Q, ldq, stream, rows_dev, nblk, blk_off);
    else launch<false>(p, rows, Q, ldq, stream, rows_dev, nblk, blk_off);
}

}  // namespace ssb
