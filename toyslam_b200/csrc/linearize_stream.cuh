// linearize_stream.cuh -- K1b for HBM-sized graphs: persistent blocks stream the LM-edge chunks.
//
// Why: the one-unit-per-block kernel (k_linearize) is block-synchronous -- load, barrier, compute,
// barrier, reduce -- so only a fraction of the resident blocks has loads in flight at any time
// (profiles/r1_c4_ncu_full_summary.csv: 2.2 TB/s, barrier + long-scoreboard stalls).  Here every
// block keeps the loads of the next chunks in flight while it computes the current one:
//   * the index / vertex arrays of chunk k+1, k+2 (lm_packed, slot_edge, slot_lm, slot_ptr, pose_ptr
//     and the unit's pose block) arrive through a 3-stage shared-memory ring filled by the TMA engine
//     (cp.async.bulk 1-D copies signalled through mbarriers, issued by one thread);
//   * the per-edge values of chunk k+1 (z_loc, information: consumed once, by the thread that owns
//     the edge) and the landmarks of its slots are prefetched into registers.
// The arithmetic and every summation order are those of unit_linearize (same device functions), so
// H, b and the slot partials are bit-identical to it.
//
// Included inside the anonymous namespace of kernels.cu.

constexpr int kLsStages = 3;
// byte offsets inside one stage (multiples of 32; cp.async.bulk needs 16-B aligned addresses and sizes)
constexpr uint32_t kLsPk = 0;                                   // lm_packed  u32 x (512 + 3 align + pad)
constexpr uint32_t kLsSe = kLsPk + 2080;                        // slot_edge  u16 x (512 + 7 + pad)
constexpr uint32_t kLsSl = kLsSe + 1056;                        // slot_lm    u32 x (512 + 3 + pad)
constexpr uint32_t kLsSp = kLsSl + 2080;                        // slot_ptr   u32 x (513 + 3 + pad)
constexpr uint32_t kLsPp = kLsSp + 2080;                        // pose_ptr   u32 x (129 + 3 + pad)
constexpr uint32_t kLsPose = kLsPp + 544;                       // pose       double4 x 128
constexpr uint32_t kLsStageBytes = kLsPose + 4096;
static_assert(kLsStageBytes % 32 == 0 && kLsPose % 32 == 0, "stage layout");

constexpr uint32_t kLsRows = 0;                                                       // 9 x kStageStride doubles
constexpr uint32_t kLsSlot2 = kLsRows + (uint32_t)align32(sizeof(double) * kStageRows * kStageStride);
constexpr uint32_t kLsRing = kLsSlot2 + sizeof(double2) * kChunkEdges;
constexpr uint32_t kLsRed = kLsRing + kLsStages * kLsStageBytes;
constexpr uint32_t kLsBar = kLsRed + 256;
constexpr uint32_t kLsSmemBytes = kLsBar + 64;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// A chunk descriptor packed into four words for the chunk loops: they carry the descriptors of four chunks at a time, and at the
// 64-register cap (2 x 512 threads per SM) the 32 registers of four plain ChunkDesc left k_pcg_stream<true> -- which keeps a few
// more values alive across the passes -- with a dozen spill reloads inside pass A's loop (pass A then took as long for half the
// chunks at 2 GPUs as for all of them at 1: profiles/r2c_pcg_phase_spread_n2.txt).  w = n | np << 10 | ns << 18 | (unit & 1) << 28:
// n, ns <= 512, np <= 128; consecutive chunks belong to the same or to consecutive units, so one bit of the unit tells them apart.
struct PDesc {
    uint32_t e0, p0, s0, w;
    __device__ __forceinline__ uint32_t n() const { return w & 1023u; }
    __device__ __forceinline__ uint32_t np() const { return (w >> 10) & 255u; }
    __device__ __forceinline__ uint32_t ns() const { return (w >> 18) & 1023u; }
    __device__ __forceinline__ uint32_t ub() const { return w >> 28; }
    __device__ __forceinline__ uint32_t e1() const { return e0 + n(); }
    __device__ __forceinline__ uint32_t p1() const { return p0 + np(); }
    __device__ __forceinline__ uint32_t s1() const { return s0 + ns(); }
};
__device__ __forceinline__ PDesc pd_pack(const ChunkDesc& ch) {
    return PDesc{ch.e0, ch.p0, ch.s0, (ch.e1 - ch.e0) | ((ch.p1 - ch.p0) << 10) | ((ch.s1 - ch.s0) << 18) | ((ch.block & 1u) << 28)};
}
__device__ __forceinline__ PDesc pd_load(const Dev& d, uint32_t c) { return pd_pack(d.chunks[c]); }

// Producer (one thread): arm the stage's barrier and start the bulk copies of chunk `ch`.
// Sources are aligned down / sizes rounded up to 16 B (every device array carries 64 B of slack).
__device__ __forceinline__ void ls_issue(const Dev& d, const PDesc& pd, unsigned char* stage, uint32_t bar, bool with_pose = true) {
    const struct { uint32_t e0, e1, p0, p1, s0, s1; } ch = {pd.e0, pd.e1(), pd.p0, pd.p1(), pd.s0, pd.s1()};
    const uint32_t n = ch.e1 - ch.e0, ns = ch.s1 - ch.s0, np = ch.p1 - ch.p0;
    const uint32_t e4 = ch.e0 & ~3u, e8 = ch.e0 & ~7u, s4 = ch.s0 & ~3u, p4 = ch.p0 & ~3u;
    const uint32_t bPk = n ? (((ch.e1 - e4) + 3u) & ~3u) * 4u : 0u;
    const uint32_t bSe = n ? (((ch.e1 - e8) + 7u) & ~7u) * 2u : 0u;
    const uint32_t bSl = ns ? (((ch.s1 - s4) + 3u) & ~3u) * 4u : 0u;
    const uint32_t bSp = (((ch.s1 + 1 - s4) + 3u) & ~3u) * 4u;
    const uint32_t bPp = (((ch.p1 + 1 - p4) + 3u) & ~3u) * 4u;
    const uint32_t bPose = with_pose ? np * 32u : 0u;
    const uint32_t s = smem_u32(stage);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the stage was last read through the generic proxy
    mbar_expect_tx(bar, bPk + bSe + bSl + bSp + bPp + bPose);
    if (n) { bulk_g2s(s + kLsPk, d.lm_packed + e4, bPk, bar); bulk_g2s(s + kLsSe, d.slot_edge + e8, bSe, bar); }
    if (ns) bulk_g2s(s + kLsSl, d.slot_lm + s4, bSl, bar);
    bulk_g2s(s + kLsSp, d.slot_ptr + s4, bSp, bar);
    bulk_g2s(s + kLsPp, d.pose_ptr + p4, bPp, bar);
    if (bPose) bulk_g2s(s + kLsPose, d.pose + ch.p0, bPose, bar);
}

constexpr int kLsThreads = 512;         // one LM edge per thread; 2 blocks (32 warps) per SM at <= 64 registers

// same association as seg_sum / list_sum (four strided partial sums, then (t0+t1)+(t2+t3)), kept as rolled
// loops: the fully unrolled variants cost more in divergent control flow than they save (ncu: 82 % of the
// issued instructions of the first streaming version were integer / branch overhead)
__device__ __forceinline__ double seg_sum_lean(const double* s, uint32_t b, uint32_t e) {
    double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    uint32_t k = b;
#pragma unroll 1
    for (; k + 3 < e; k += 4) { t0 += s[k]; t1 += s[k + 1]; t2 += s[k + 2]; t3 += s[k + 3]; }
#pragma unroll 1
    for (; k < e; ++k) t0 += s[k];
    return (t0 + t1) + (t2 + t3);
}
__device__ __forceinline__ double list_sum_lean(const double* s, const uint16_t* idx, uint32_t b, uint32_t e) {
    double t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    uint32_t k = b;
#pragma unroll 1
    for (; k + 3 < e; k += 4) { t0 += s[idx[k]]; t1 += s[idx[k + 1]]; t2 += s[idx[k + 2]]; t3 += s[idx[k + 3]]; }
#pragma unroll 1
    for (; k < e; ++k) t0 += s[idx[k]];
    return (t0 + t1) + (t2 + t3);
}

// reduction item (pose, row) -> where its value goes: Hpp (00 01 02 11 12 22) or b_p, and which entry of
// pose_od (the ODOM / prior part prepared by k_odom) is added; b_2 = od - sum (b -= A^T W e)
__device__ __forceinline__ int ls_od_index(uint32_t row) { return row == 0 ? 0 : row == 2 ? 1 : row == 5 ? 2 : row >= 6 ? (int)row - 3 : -1; }
__device__ __forceinline__ void ls_store(const Dev& d, uint32_t p, uint32_t row, double sum, double od) {
    if (row < 6) { const int hi = row == 2 ? 3 : row == 3 ? 2 : (int)row; d.hpp[6ull * p + hi] = sum + od; }
    else if (row < 8) d.bp[3ull * p + (row - 6)] = sum + od;
    else d.bp[3ull * p + 2] = od - sum;
}

__global__ void __launch_bounds__(kLsThreads, 2) k_linearize_stream(const Dev d) {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    double* rows = reinterpret_cast<double*>(smem_raw + kLsRows);
    double2* slot2 = reinterpret_cast<double2*>(smem_raw + kLsSlot2);
    unsigned char* ring = smem_raw + kLsRing;
    double* red = reinterpret_cast<double*>(smem_raw + kLsRed);
    const uint32_t bar0 = smem_u32(smem_raw + kLsBar);
    const uint32_t tid = threadIdx.x;

    // this block's units: a contiguous range (units carry <= kChunkEdges edges each, so ranges are balanced)
    const uint32_t u_lo = (uint32_t)((uint64_t)d.nBlocks * blockIdx.x / gridDim.x);
    const uint32_t u_hi = (uint32_t)((uint64_t)d.nBlocks * (blockIdx.x + 1) / gridDim.x);
    const uint32_t cb = d.blk_chunk_ptr[u_lo], ce = d.blk_chunk_ptr[u_hi];

    if (tid == 0) {
        for (int s = 0; s < kLsStages; ++s) mbar_init(bar0 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (cb >= ce) { if (tid == 0) d.part[blockIdx.x] = 0; return; }

    // chunk descriptors, read by every thread (uniform loads) up to three chunks ahead
    // (descriptors beyond the block's range are never used: loads clamped, uses guarded by the chunk index)
    PDesc cur = pd_load(d, cb);
    PDesc nx1 = pd_load(d, min(cb + 1, ce - 1));
    PDesc nx2 = pd_load(d, min(cb + 2, ce - 1));
    if (tid == 0) {
        ls_issue(d, cur, ring, bar0);
        if (cb + 1 < ce) ls_issue(d, nx1, ring + kLsStageBytes, bar0 + 8);
        if (cb + 2 < ce) ls_issue(d, nx2, ring + 2 * kLsStageBytes, bar0 + 16);
    }
    // per-edge values of the first chunk (afterwards prefetched one chunk ahead)
    double2 z0 = {}, w0 = {};
    if (tid < cur.n()) { z0 = d.le_z[cur.e0 + tid]; w0 = d.le_w[cur.e0 + tid]; }

    double heavy = 0;            // a pose with > kChunkEdges edges spans several chunks: its nine sums live in threads 0..8
    double rho_sum = 0;

    for (uint32_t c = cb; c < ce; ++c) {
        const uint32_t k = c - cb, sidx = k % kLsStages;
        unsigned char* st = ring + sidx * kLsStageBytes;
        const uint32_t n = cur.n(), ns = cur.ns(), np = cur.np(), e1 = cur.e0 + n;
        const bool has1 = c + 1 < ce;
        const bool last = !has1 || nx1.ub() != cur.ub();          // the block's final chunk ends its unit
        const PDesc nx3 = pd_load(d, min(c + 3, ce - 1));         // consumed at the end of this iteration
        mbar_wait(bar0 + 8 * sidx, (k / kLsStages) & 1u);
        const uint32_t* pk = reinterpret_cast<const uint32_t*>(st + kLsPk) + (cur.e0 & 3u);
        const uint16_t* sedge = reinterpret_cast<const uint16_t*>(st + kLsSe) + (cur.e0 & 7u);
        const uint32_t* slm = reinterpret_cast<const uint32_t*>(st + kLsSl) + (cur.s0 & 3u);
        const uint32_t* sptr = reinterpret_cast<const uint32_t*>(st + kLsSp) + (cur.s0 & 3u);    // absolute edge positions
        const uint32_t* pptr = reinterpret_cast<const uint32_t*>(st + kLsPp) + (cur.p0 & 3u);
        const double4* pose = reinterpret_cast<const double4*>(st + kLsPose);

        if (k == 0) {        // first chunk of the block: its landmarks have not been prefetched
            if (tid < ns) slot2[tid] = d.lm[slm[tid]];
            __syncthreads();
        }
        // ---- prefetch for the next chunk: per-edge values and slot landmarks into registers -------------------
        double2 nz0 = {}, nw0 = {}, lmv0 = {};
        const uint32_t ns_next = has1 ? nx1.ns() : 0u;
        if (has1 && tid < nx1.n()) { nz0 = d.le_z[nx1.e0 + tid]; nw0 = d.le_w[nx1.e0 + tid]; }
        if (c + 1 < ce) {
            const uint32_t s2 = (k + 1) % kLsStages;
            mbar_wait(bar0 + 8 * s2, ((k + 1) / kLsStages) & 1u);          // issued two iterations ago
            const uint32_t* slm2 = reinterpret_cast<const uint32_t*>(ring + s2 * kLsStageBytes + kLsSl) + (nx1.s0 & 3u);
            if (tid < ns_next) lmv0 = d.lm[slm2[tid]];
        }
        // the ODOM / prior part of this unit's poses for this thread's first reduction item
        const uint32_t nitems = np * kStageRows;
        double od0 = 0;
        if (last && tid < nitems) { const int oi = ls_od_index(tid / np); if (oi >= 0) od0 = d.pose_od[6ull * (cur.p0 + tid % np) + oi]; }

        // ---- per-edge linearisation, products staged in `rows`, Hpl written once ------------------------------
        if (tid < n) { const uint32_t q = pk[tid];
            rho_sum += lm_edge(pose[q >> 16], slot2[q & 0xffffu], z0, w0, hpl_row(d, cur.e0 + tid), hpl_stride(d), rows + tid); }
        __syncthreads();

        // ---- pose side: one (pose, row) item per thread sums its row segment ----------------------------------
        const bool multi = np == 1 && pptr[1] - pptr[0] > kChunkEdges;      // only a single heavy pose is ever split
#pragma unroll 1
        for (uint32_t item = tid; item < nitems; item += kLsThreads) {
            // item -> (row, pose) with the POSE varying fastest: the lanes of a warp then read one staged row at offsets that grow by
            // the poses' edge counts (19 at C4: 3 mod 16 double-banks, conflict-free) instead of nine rows of three or four poses
            // whose offsets collide (2.6x the conflict-free wavefronts, profiles/r1_lin_stream2_c4_ncu_summary.csv)
            const uint32_t row = item / np, pl = item - row * np;
            const uint32_t b = max(pptr[pl], cur.e0) - cur.e0, e = min(pptr[pl + 1], e1) - cur.e0;
            double sum = seg_sum_lean(rows + row * kStageStride, b, e);
            if (multi) { heavy += sum; sum = heavy; }             // np == 1: item == tid < 9; same order as acc += seg_sum
            if (last) {
                double od = od0;
                if (item >= kLsThreads) { const int oi = ls_od_index(row); od = oi >= 0 ? d.pose_od[6ull * (cur.p0 + pl) + oi] : 0.0; }
                ls_store(d, cur.p0 + pl, row, sum, od);
            }
        }
        if (last) heavy = 0;
        // ---- landmark side: one partial per slot and component, edges of the slot in list order ----------------
        // (items are dealt from the top thread down: the pose items above occupy the low threads, so both run at once)
#pragma unroll 1
        for (uint32_t item = kLsThreads - 1 - tid; item < ns * 5; item += kLsThreads) {
            const uint32_t sl = item / 5, comp = item % 5;
            const double* s = rows + (comp < 3 ? comp : comp + 3) * kStageStride;
            d.lm_part[5ull * (cur.s0 + sl) + comp] = list_sum_lean(s, sedge, sptr[sl] - cur.e0, sptr[sl + 1] - cur.e0);
        }
        // slot2 is no longer read by this chunk: hand it to the next one
        if (tid < ns_next) slot2[tid] = lmv0;
        __syncthreads();
        // stage `sidx` is free: refill it with the chunk three ahead
        if (tid == 0 && c + 3 < ce) ls_issue(d, nx3, st, bar0 + 8 * sidx);
        cur = nx1; nx1 = nx2; nx2 = nx3;
        z0 = nz0; w0 = nw0;
    }
    const double t = block_sum(rho_sum, red);
    if (tid == 0) d.part[blockIdx.x] = t;
}
