// The epoch replay of lo_robin_block.cuh for key sets that do not fit one block's shared memory (update sets of more than
// 16 384 points: a 128-beam scan has ~180 k): the same phases, run by the whole cooperative grid on 32-bit arrays in global
// memory (L2 resident), with grid-wide barriers where the block version has block barriers. It continues where the block
// version stops: block 0 replays the first 2^12 elements in shared memory (epochs 0..12, where a grid-wide barrier per
// phase would cost more than the work), the grid takes the epochs from there — two for a VLP16 update set, six for the
// 180 k keys of a 128-beam scan.
// See lo_robin_block.cuh for the algorithm; the phase names are the same.
#pragma once

#include <cooperative_groups.h>
#include <limits.h>
#include <stdint.h>

#include "lo_robin_block.cuh"

namespace lo {

namespace cg = cooperative_groups;

struct RobinGridBufs {        // global scratch for n keys, C = 2 pow2ceil(n) buckets
    uint32_t* start;          // [C]   counts -> exclusive starts (bit 31: the bucket heads a cluster) -> ends after the placement
    uint32_t* seq;            // [n]   sequence indices grouped by bucket (sorted position t)
    uint32_t* stash;          // [n]   ideal bucket of sequence element s; after R3: the occupied cells in position order
    uint32_t* cells;          // [n]   cluster label (sorted position of the cluster's first element; there: bit 31 | members)
    uint32_t* prev[2];        // [n]   iteration order of the previous epoch (element ids), double-buffered
    uint32_t* atime;          // [n]   ideal cell of a cluster's k-th insertion, at sorted position t0 + k
    uint32_t* etime;          // [n]   the empty cell it filled
    uint32_t* rank;           // [n]   rank of the member at sorted position t inside its cluster
    uint32_t* part;           // [2][2 * blocks] per-block totals of the grid-wide scans
    uint32_t* flags;          // [0] cells past the last bucket (wrap), [1] refused
};
constexpr uint32_t kRobinGridMaxBlocks = 1024;
__host__ __device__ inline size_t robin_grid_words(uint32_t n) {   // 32-bit words of scratch for n keys (the caller carves them)
    uint32_t c = 2;
    while (c < 2 * n) c <<= 1;
    return static_cast<size_t>(c) + 8 * static_cast<size_t>(n) + 4 * kRobinGridMaxBlocks + 8;
}
__host__ __device__ inline RobinGridBufs robin_grid_carve(uint32_t* base, uint32_t n) {
    uint32_t c = 2;
    while (c < 2 * n) c <<= 1;
    RobinGridBufs b;
    b.start = base; base += c;
    b.seq = base; base += n;
    b.stash = base; base += n;
    b.cells = base; base += n;
    b.prev[0] = base; base += n;
    b.prev[1] = base; base += n;
    b.atime = base; base += n;
    b.etime = base; base += n;
    b.rank = base; base += n;
    b.part = base; base += 4 * kRobinGridMaxBlocks;
    b.flags = base;
    return b;
}

// exclusive prefix sum and exclusive prefix maximum of one value per thread over the whole grid (threads in global index
// order), and the grid totals. One grid-wide barrier; `slot` alternates between consecutive calls.
__device__ inline void robin_grid_scan(cg::grid_group& grid, uint32_t v, int g, uint32_t& v_before, int& g_before, uint32_t& v_all, int& g_all,
                                       uint32_t* part, int slot) {
    uint32_t vb, va;
    int gb, ga;
    robin_block_scan(v, g, vb, gb, va, ga, slot);
    uint32_t* mine = part + static_cast<size_t>(slot) * 2 * kRobinGridMaxBlocks;
    if (threadIdx.x == 0) {
        mine[2 * blockIdx.x] = va;
        mine[2 * blockIdx.x + 1] = static_cast<uint32_t>(ga);
    }
    grid.sync();
    __shared__ uint32_t s_sum[2][2];
    __shared__ int s_max[2][2];
    if (threadIdx.x == 0) {
        s_sum[slot][0] = s_sum[slot][1] = 0u;
        s_max[slot][0] = s_max[slot][1] = INT_MIN;
    }
    __syncthreads();
    uint32_t before_v = 0, all_v = 0;
    int before_g = INT_MIN, all_g = INT_MIN;
    for (uint32_t k = threadIdx.x; k < gridDim.x; k += blockDim.x) {
        const uint32_t pv = __ldcg(mine + 2 * k);
        const int pg = static_cast<int>(__ldcg(mine + 2 * k + 1));
        all_v += pv;
        all_g = max(all_g, pg);
        if (k < blockIdx.x) {
            before_v += pv;
            before_g = max(before_g, pg);
        }
    }
    for (int off = 16; off > 0; off >>= 1) {
        before_v += __shfl_xor_sync(0xFFFFFFFFu, before_v, off);
        all_v += __shfl_xor_sync(0xFFFFFFFFu, all_v, off);
        before_g = max(before_g, __shfl_xor_sync(0xFFFFFFFFu, before_g, off));
        all_g = max(all_g, __shfl_xor_sync(0xFFFFFFFFu, all_g, off));
    }
    if ((threadIdx.x & 31) == 0 && threadIdx.x < gridDim.x + 31) {
        atomicAdd(&s_sum[slot][0], before_v);
        atomicAdd(&s_sum[slot][1], all_v);
        atomicMax(&s_max[slot][0], before_g);
        atomicMax(&s_max[slot][1], all_g);
    }
    __syncthreads();
    v_before = s_sum[slot][0] + vb;
    g_before = max(s_max[slot][0], gb);
    v_all = s_sum[slot][1];
    g_all = s_max[slot][1];
}

// uh[k]: 22-bit hash of the k-th key in insertion order, k < n. out[d] = ids[element at iteration position d] (ids == nullptr:
// the element index itself). Epochs j0, j0 + 1, ... are run; B.prev[0][0 .. lo0) must hold the iteration order after epoch
// j0 - 1 (lo0 = 2^(j0-1) elements; j0 = 0, lo0 = 0: from scratch). Every thread of the cooperative grid calls.
// Returns false (in every thread; `out` untouched) if a cluster exceeded kRobinBlockMaxCluster.
__device__ inline bool robin_order_grid(cg::grid_group& grid, const uint32_t* __restrict__ uh, const uint32_t* __restrict__ ids, uint32_t n,
                                        const RobinGridBufs& B, uint32_t* __restrict__ out, uint32_t j0, uint32_t lo0) {
    const uint32_t gt = blockIdx.x * blockDim.x + threadIdx.x, T = gridDim.x * blockDim.x;
    constexpr uint32_t kHead = 0x80000000u, kVal = 0x7FFFFFFFu;
    uint32_t lo = lo0;
    int cur = 0;
    if (gt == 0) B.flags[1] = 0u;
    for (uint32_t j = j0; lo < n; j++) {
        const uint32_t C = 2u << j, mask = C - 1;
        const uint32_t hi = j == 0 ? 1u : min(n, 1u << j);
        const uint32_t ne = hi;
        const uint32_t* prev = B.prev[cur];
        uint32_t* next = B.prev[cur ^ 1];
        auto at = [&](uint32_t x) { return x >= ne ? x - ne : x; };
        // E1
        for (uint32_t b = gt; b < C; b += T) B.start[b] = 0u;
        grid.sync();
        // E2
        for (uint32_t s = gt; s < ne; s += T) {
            const uint32_t e = s < lo ? __ldcg(prev + s) : s;
            const uint32_t b = __ldg(uh + e) & mask;
            B.stash[s] = b;
            atomicAdd(B.start + b, 1u);
        }
        grid.sync();
        // E3
        const uint32_t per = (C + T - 1) / T;
        const uint32_t b_lo = min(C, gt * per), b_hi = min(C, b_lo + per);
        uint32_t sum = 0;
        int m = INT_MIN;
        for (uint32_t b = b_lo; b < b_hi; b++) {
            const uint32_t c = __ldcg(B.start + b);
            if (c) m = max(m, static_cast<int>(b) - static_cast<int>(sum));
            sum += c;
        }
        uint32_t base, unused_all;
        int unused_g, unused_ga;
        robin_grid_scan(grid, sum, INT_MIN, base, unused_g, unused_all, unused_ga, B.part, 0);
        uint32_t unused_vb, unused_va;
        int g_before, g_all;
        robin_grid_scan(grid, 0u, m == INT_MIN ? INT_MIN : m - static_cast<int>(base), unused_vb, g_before, unused_va, g_all, B.part, 1);
        {
            uint32_t run = base;
            int G = g_before;
            for (uint32_t b = b_lo; b < b_hi; b++) {
                const uint32_t c = __ldcg(B.start + b);
                uint32_t v = run;
                if (c) {
                    const int g = static_cast<int>(b) - static_cast<int>(run);
                    if (g > G) v |= kHead;
                    G = max(G, g);
                }
                B.start[b] = v;
                run += c;
            }
        }
        const int over = static_cast<int>(ne) + g_all - static_cast<int>(C);
        const uint32_t w = over > 0 ? static_cast<uint32_t>(over) : 0u;   // identical in every thread
        grid.sync();
        // E4
        if (gt == 0 && w) {
            for (uint32_t o = 0; o < C; o++) {
                const uint32_t s0 = __ldcg(B.start + o) & kVal, s1 = o + 1 < C ? (__ldcg(B.start + o + 1) & kVal) : ne;
                if (s1 == s0) continue;
                if (static_cast<int>(o) - static_cast<int>(s0) > static_cast<int>(w)) break;
                B.start[o] = s0;
            }
        }
        grid.sync();
        // E5
        for (uint32_t s = gt; s < ne; s += T) {
            const uint32_t b = __ldcg(B.stash + s);
            const uint32_t t = atomicAdd(B.start + b, 1u) & kVal;
            B.seq[t] = s;
        }
        grid.sync();
        // A
        {
            const uint32_t q = (ne + T - 1) / T;
            if (q > 64) {   // cannot happen on a B200 grid (n <= 2^19 over 75 776 threads); refuse rather than overflow the mask
                if (gt == 0) B.flags[1] = 1u;
            }
            const uint32_t t_lo = min(ne, gt * q), t_hi = q > 64 ? t_lo : min(ne, t_lo + q);
            unsigned long long firsts = 0;
            int last_first = INT_MIN;
            for (uint32_t t = t_lo; t < t_hi; t++) {
                const uint32_t b = __ldcg(B.stash + __ldcg(B.seq + t));
                if ((__ldcg(B.start + b) & kHead) && t == (b ? (__ldcg(B.start + b - 1) & kVal) : 0u)) {
                    firsts |= 1ull << (t - t_lo);
                    last_first = static_cast<int>(t);
                }
            }
            uint32_t unused_b, unused_a;
            int before, overall;
            robin_grid_scan(grid, 0u, last_first, unused_b, before, unused_a, overall, B.part, 0);
            int label = before == INT_MIN ? overall : before;
            for (uint32_t t = t_lo; t < t_hi; t++) {
                if ((firsts >> (t - t_lo)) & 1ull) label = static_cast<int>(t);
                B.cells[t] = static_cast<uint32_t>(label);
            }
        }
        grid.sync();
        // R2
        for (uint32_t t = gt; t < ne; t += T) {
            const uint32_t v = __ldcg(B.cells + t);
            if (!(v & kHead) && v != t) continue;
            const uint32_t t_last = t ? t - 1 : ne - 1, pv = __ldcg(B.cells + t_last);
            const uint32_t t_prev = (pv & kHead) ? t_last : pv;
            const uint32_t len = t_prev == t ? ne : (t > t_prev ? t - t_prev : t + ne - t_prev);
            B.cells[t_prev] = kHead | len;
            if (len > kRobinBlockMaxCluster) B.flags[1] = 1u;
        }
        grid.sync();
        if (__ldcg(B.flags + 1)) return false;
        auto cluster_of = [&](uint32_t t, uint32_t& t0, uint32_t& L) {
            const uint32_t v = __ldcg(B.cells + t);
            t0 = (v & kHead) ? t : v;
            L = (v & kHead) ? (v & kVal) : (__ldcg(B.cells + t0) & kVal);
        };
        // R3
        for (uint32_t t = gt; t < ne; t += T) {
            uint32_t t0, L;
            cluster_of(t, t0, L);
            const uint32_t mine = __ldcg(B.seq + t);
            const uint32_t a = (__ldcg(B.stash + mine) - __ldcg(B.stash + __ldcg(B.seq + t0))) & mask;
            uint32_t r = 0;
            for (uint32_t i = 0; i < L; i++) r += __ldcg(B.seq + at(t0 + i)) < mine ? 1u : 0u;
            B.rank[t] = r;
            B.atime[at(t0 + r)] = a;
        }
        grid.sync();
        // P
        for (uint32_t t0 = gt; t0 < ne; t0 += T) {
            const uint32_t v = __ldcg(B.cells + t0);
            if (!(v & kHead)) continue;
            const uint32_t L = v & kVal;
            if (L <= 64) {
                unsigned long long occupied = 0;
                for (uint32_t k = 0; k < L; k++) {
                    const uint32_t u = at(t0 + k), a = __ldcg(B.atime + u);
                    const uint32_t e = a + static_cast<uint32_t>(__ffsll(static_cast<long long>(~occupied >> a))) - 1u;
                    occupied |= 1ull << e;
                    B.etime[u] = e;
                }
            } else {   // a long cluster: occupancy of cell i in bit 31 of atime[t0 + i]; this thread alone touches these words
                for (uint32_t k = 0; k < L; k++) {
                    const uint32_t u = at(t0 + k);
                    uint32_t e = __ldcg(B.atime + u) & kVal;
                    while (__ldcg(B.atime + at(t0 + e)) & kHead) e++;
                    B.atime[at(t0 + e)] = __ldcg(B.atime + at(t0 + e)) | kHead;
                    B.etime[u] = e;
                }
            }
        }
        grid.sync();
        // W
        for (uint32_t t = gt; t < ne; t += T) {
            uint32_t t0, L;
            cluster_of(t, t0, L);
            const uint32_t r = B.rank[t], a = __ldcg(B.atime + at(t0 + r)) & kVal;
            uint32_t before = 0, idx = 0;
            for (uint32_t i = 0; i < r; i++) {
                const uint32_t a2 = __ldcg(B.atime + at(t0 + i)) & kVal;
                before += a2 < a ? 1u : 0u;
                idx += a2 == a ? 1u : 0u;
            }
            uint32_t size = idx + 1u;
            for (uint32_t i = r + 1u; i < L; i++) {
                const uint32_t u = at(t0 + i), a2 = __ldcg(B.atime + u) & kVal, e2 = __ldcg(B.etime + u);
                before += a2 < a ? 1u : 0u;
                if (a2 == a) size++;
                else if (a2 < a && a < e2) idx = idx ? idx - 1u : size - 1u;
            }
            B.stash[at(t0 + before + idx)] = __ldcg(B.seq + t);
        }
        grid.sync();
        // E7
        const bool last = hi == n;
        for (uint32_t t = gt; t < ne; t += T) {
            uint32_t d = t + w;
            if (d >= ne) d -= ne;
            const uint32_t s = __ldcg(B.stash + t);
            const uint32_t e = s < lo ? __ldcg(prev + s) : s;
            if (last) out[d] = ids ? __ldg(ids + e) : e;
            else next[d] = e;
        }
        grid.sync();
        cur ^= 1;
        lo = hi;
    }
    return true;
}

// The order for any n: block 0 replays the first 2^handover elements in its shared memory (`smem`: kRobinBlockSmemBytes, needed in
// block 0 only; epochs 0 .. handover, where a grid-wide barrier per phase would cost more than the work), the grid takes the
// epochs after that. handover <= 14; 0: every epoch on the grid (tests). Every thread of the cooperative grid calls; returns
// false (uniformly) if the key set was refused.
constexpr uint32_t kRobinGridMax = 1u << 19;   // keys: what the per-scan kernels take as a scan
constexpr uint32_t kRobinHandover = 12;   // measured: 13 000 keys take 0.27 ms with 2^12 against 0.33 ms with 2^14 (all in the block) and 0.31 ms with 2^10
__device__ inline bool robin_order_any(cg::grid_group& grid, const uint32_t* __restrict__ uh, const uint32_t* __restrict__ ids, uint32_t n,
                                       unsigned char* smem, const RobinGridBufs& B, uint32_t* __restrict__ out, uint32_t handover) {
    const uint32_t first = 1u << handover;
    if (handover == 0) return robin_order_grid(grid, uh, ids, n, B, out, 0u, 0u);
    if (n <= first) {
        if (blockIdx.x == 0) {
            const bool ok = robin_order_block(uh, ids, n, smem, out);
            if (threadIdx.x == 0) B.flags[1] = ok ? 0u : 1u;
        }
        grid.sync();
        return __ldcg(B.flags + 1) == 0u;
    }
    if (blockIdx.x == 0) {   // epochs 0 .. handover: the iteration order of the first 2^handover elements, as element indices
        const bool ok = robin_order_block(uh, nullptr, first, smem, B.prev[0]);
        if (threadIdx.x == 0) B.flags[1] = ok ? 0u : 1u;
    }
    grid.sync();
    if (__ldcg(B.flags + 1)) return false;
    return robin_order_grid(grid, uh, ids, n, B, out, handover + 1u, first);
}

}  // namespace lo
