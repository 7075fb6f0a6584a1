// The exact iteration order of the reference's tsl::robin_map (third-party/robin-map, robin_hash.h) after n insertions of
// distinct keys, computed by ONE thread block in shared memory — the order keyframe_downsampler.getCloud() hands the update
// cloud to keyframe_.addCloud in (src/lidar_odometry.cpp:38,42,69), for the per-scan kernel.
//
// Why not a counting sort: inside a group of elements with the same ideal bucket the order depends on the whole history —
// an insertion into an earlier group of the same cluster moves the first element of every following group to that group's
// end (robin_hash.h:1283-1317: swap only if strictly richer), and every growth re-inserts everything in the old iteration
// order (:1319-1341). That order decides which point a full key-frame voxel drops (voxel_grid.h:82-91): approximating it
// (ties in arrival order) left 200 of 1.3 M key-frame points different after 250 scans of a stream, and the poses followed.
//
// How it parallelises. The map lives through EPOCHS: inserting element k (0-based) happens in a table of C = 2 pow2ceil(k+1)
// buckets (growth x2 from 0 at load 0.5, robin_hash.h:1379-1403); epoch j has C = 2^(j+1) and starts by re-inserting the
// 2^(j-1) elements present in the previous epoch's iteration order, followed by the arrivals 2^(j-1) .. 2^j - 1. Within an
// epoch the set of occupied cells does not depend on the insertion order (linear probing), so the final CLUSTERS (maximal
// runs of occupied cells) are known up front, nothing ever enters or leaves a final cluster's cells, and the history of a
// cluster involves its own members only. So per epoch: counting sort of the sequence by ideal bucket, cluster heads from a
// prefix maximum (element t of the sorted order sits at t + max_{u<=t}(b_u - u); a new cluster starts where that maximum
// rises), and then the history of every cluster WITHOUT replaying it cell by cell:
//   * robin_hash.h:1283-1317 swaps only with a strictly richer resident, and the cells of a run are sorted by ideal bucket,
//     so the insertion of x (ideal cell a) does exactly this: x goes to the end of its own group (the elements with ideal
//     cell a), and every following group of the run up to the first empty cell e has its first element carried to its
//     end (a rotation by one; the group moves one cell up). e is the first empty cell >= a — plain linear probing.
//   * hence, per cluster: (1) the members in sequence order (rank by counting, all members in parallel); (2) e of every
//     insertion, one thread per cluster with an occupancy bitmap in registers; (3) every member on its own: start at the
//     end of its group, then walk the LATER insertions (a', e') of the cluster: one of its own group -> the group grew;
//     a' < a < e' -> the group was rotated: index 0 goes to the end, the others move down. Its final cell is the group's
//     first cell + that index. O(cluster length) independent loads per member instead of a serial replay per cluster.
// The iteration order of the epoch's table (the clusters' cells in position order; the w cells that ran past the last
// bucket come first) is the next epoch's re-insertion sequence, and the last one is the answer. Identical to the
// sequential replay of lo_robin.cu, which tests/test_gpu_robin.py pins to the real container.
#pragma once

#include <limits.h>
#include <stdint.h>

namespace lo {

constexpr uint32_t kRobinBlockMax = 16384;          // elements the shared-memory replay takes (a VLP16 update set is 12-15 k)
constexpr uint32_t kRobinBlockBuckets = 2 * kRobinBlockMax;
// A cluster longer than this is refused (the function returns false): step (2) is serial per cluster (and quadratic beyond
// 64 cells), and past 8192 cells tsl's probe-distance limit (robin_hash.h:327) adds growth steps this replay does not model.
// Hash-random keys at load <= 0.5 give clusters of tens of cells; the sequential replay of lo_robin.cu takes anything.
constexpr uint32_t kRobinBlockMaxCluster = 2048;
// uint16 arrays: start[C] | seq[n] | stash[n] | cells[n] | prev[n / 2]
constexpr size_t kRobinBlockSmemBytes = 2 * (kRobinBlockBuckets + 3 * kRobinBlockMax + kRobinBlockMax / 2);

struct RobinBlockSmem {
    uint16_t* start;   // [C]  counts -> exclusive starts (bit 15: the bucket heads a cluster) -> ends after the placement;
                       //      dead after the labels: [0, n) = ideal cell of a cluster's k-th insertion, [n, 2n) = the cell it filled
    uint16_t* seq;     // [n]  sequence indices grouped by bucket (sorted position t); E7: scratch
    uint16_t* stash;   // [n]  ideal bucket of sequence element s; dead after R3: the occupied cells in position order from the
                       //      first non-wrapped one: sequence index of the resident
    uint16_t* cells;   // [n]  cluster label (sorted position of the cluster's first element; there: bit 15 | members)
    uint16_t* prev;    // [n/2] previous epoch's iteration order (element ids)
};
__device__ inline RobinBlockSmem robin_block_carve(unsigned char* base) {
    RobinBlockSmem s;
    s.start = reinterpret_cast<uint16_t*>(base);
    s.seq = s.start + kRobinBlockBuckets;
    s.stash = s.seq + kRobinBlockMax;
    s.cells = s.stash + kRobinBlockMax;
    s.prev = s.cells + kRobinBlockMax;
    return s;
}

// exclusive prefix sum and exclusive prefix maximum of one value per thread, and the block totals; blockDim.x <= 1024.
// One block barrier per call: every warp scans the per-warp totals itself; `slot` alternates between consecutive calls
// (two sets of partials, so a warp still reading one call's partials is not overrun by the next call).
__device__ inline void robin_block_scan(uint32_t v, int g, uint32_t& v_before, int& g_before, uint32_t& v_all, int& g_all, int slot) {
    __shared__ uint32_t ws[2][32];
    __shared__ int wg[2][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    uint32_t vi = v;
    int gi = g;
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t uv = __shfl_up_sync(0xFFFFFFFFu, vi, off);
        const int ug = __shfl_up_sync(0xFFFFFFFFu, gi, off);
        if (lane >= off) {
            vi += uv;
            gi = max(gi, ug);
        }
    }
    if (lane == 31) {
        ws[slot][warp] = vi;
        wg[slot][warp] = gi;
    }
    __syncthreads();
    uint32_t pv = lane < nwarps ? ws[slot][lane] : 0u;   // inclusive scan of the warp totals, in every warp
    int pg = lane < nwarps ? wg[slot][lane] : INT_MIN;
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t uv = __shfl_up_sync(0xFFFFFFFFu, pv, off);
        const int ug = __shfl_up_sync(0xFFFFFFFFu, pg, off);
        if (lane >= off) {
            pv += uv;
            pg = max(pg, ug);
        }
    }
    v_all = __shfl_sync(0xFFFFFFFFu, pv, 31);
    g_all = __shfl_sync(0xFFFFFFFFu, pg, 31);
    const uint32_t vb = __shfl_sync(0xFFFFFFFFu, pv, max(warp - 1, 0));
    const int gb = __shfl_sync(0xFFFFFFFFu, pg, max(warp - 1, 0));
    const uint32_t v_excl_in_warp = __shfl_up_sync(0xFFFFFFFFu, vi, 1);
    const int g_excl_in_warp = __shfl_up_sync(0xFFFFFFFFu, gi, 1);
    const uint32_t warp_v = warp ? vb : 0u;
    const int warp_g = warp ? gb : INT_MIN;
    v_before = warp_v + (lane ? v_excl_in_warp : 0u);
    g_before = lane ? max(warp_g, g_excl_in_warp) : warp_g;
}

// elements (16 bit) per thread for a contiguous-chunk pass: enough to cover `count` with `nthreads` threads, an even number
// whose half is odd, so that the threads of a warp start in different shared-memory banks
__device__ __forceinline__ uint32_t robin_block_chunk(uint32_t count, uint32_t nthreads) {
    return 2u * (((count + 2u * nthreads - 1u) / (2u * nthreads)) | 1u);
}

// uh[k]: the reference's 22-bit hash of the k-th key in insertion (arrival) order, k < n <= kRobinBlockMax (global memory).
// ids[k]: what to emit for element k (nullptr: k itself). out[d] = ids[element at iteration position d]. All threads of ONE block of >= 512
// threads call. Returns false (in every thread; `out` is then unspecified) if a cluster exceeded kRobinBlockMaxCluster.
__device__ inline bool robin_order_block(const uint32_t* __restrict__ uh, const uint32_t* __restrict__ ids, uint32_t n, unsigned char* smem,
                                         uint32_t* __restrict__ out, long long* prof = nullptr) {
    // prof (developer aid, LO_ROBIN_PROF=1 with lo_robin_iteration_order): [phase] cycles summed over epochs, [32 + j] per epoch
    long long prof_last = prof ? clock64() : 0;
    auto mark = [&](int slot) {
        if (prof && threadIdx.x == 0) {
            const long long c = clock64();
            prof[slot] += c - prof_last;
            prof_last = c;
        }
    };
    const RobinBlockSmem sm = robin_block_carve(smem);
    __shared__ uint32_t wrap_s;
    __shared__ int refused_s;
    if (threadIdx.x == 0) refused_s = blockDim.x < 512 ? 1 : 0;
    __syncthreads();
    if (refused_s) return false;
    const uint32_t tid = threadIdx.x, nthreads = blockDim.x;
    constexpr uint16_t kHead = 0x8000u, kVal = 0x7FFFu;
    constexpr int kMaxOwn = kRobinBlockMax / 512;   // members per thread
    uint16_t l_r[kMaxOwn];   // rank of my k-th member inside its cluster (the only state a thread carries between phases)
    uint32_t lo = 0;   // elements present before this epoch's arrivals = length of `prev`
    for (uint32_t j = 0; lo < n; j++) {
        const uint32_t C = 2u << j, mask = C - 1;
        const uint32_t hi = j == 0 ? 1u : min(n, 1u << j);   // arrivals of this epoch: element ids lo .. hi - 1
        const uint32_t ne = hi;                               // sequence S = prev[0 .. lo) ++ [lo .. hi): S[s] = s for s >= lo
        uint32_t* packed = reinterpret_cast<uint32_t*>(sm.start);
        auto at = [&](uint32_t x) { return x >= ne ? x - ne : x; };   // sorted positions are cyclic (the cluster that wraps)
        // E1: clear the counts
        for (uint32_t w = tid; w < C / 2; w += nthreads) packed[w] = 0;
        __syncthreads();
        mark(1);
        // E2: histogram of the ideal buckets (two 16-bit counters per word; a count is at most ne < 2^15)
        for (uint32_t s = tid; s < ne; s += nthreads) {
            const uint32_t e = s < lo ? sm.prev[s] : s;
            const uint32_t b = __ldg(uh + e) & mask;
            sm.stash[s] = static_cast<uint16_t>(b);
            atomicAdd(packed + (b >> 1), (b & 1u) ? 65536u : 1u);
        }
        __syncthreads();
        mark(2);
        // E3: counts -> exclusive starts, cluster heads, overflow past the last bucket. Thread t owns `per` consecutive buckets.
        const uint32_t per = robin_block_chunk(C, nthreads);
        const uint32_t b_lo = min(C, tid * per), b_hi = min(C, b_lo + per);
        uint32_t sum = 0;
        int m = INT_MIN;   // max over my non-empty buckets of b - (elements of my earlier buckets)
        for (uint32_t b = b_lo; b < b_hi; b++) {
            const uint32_t c = sm.start[b];
            if (c) m = max(m, static_cast<int>(b) - static_cast<int>(sum));
            sum += c;
        }
        uint32_t base, unused_all;
        int unused_g, unused_ga;
        robin_block_scan(sum, INT_MIN, base, unused_g, unused_all, unused_ga, 0);
        uint32_t unused_vb, unused_va;
        int g_before, g_all;   // g_b = b - start_b: where bucket b's first element would sit relative to its sorted index
        robin_block_scan(0u, m == INT_MIN ? INT_MIN : m - static_cast<int>(base), unused_vb, g_before, unused_va, g_all, 1);
        {
            uint32_t run = base;
            int G = g_before;
            for (uint32_t b = b_lo; b < b_hi; b++) {
                const uint32_t c = sm.start[b];
                uint16_t v = static_cast<uint16_t>(run);
                if (c) {
                    const int g = static_cast<int>(b) - static_cast<int>(run);
                    if (g > G) v |= kHead;   // the bucket's first element sits at its ideal cell: nothing before it reaches this far
                    G = max(G, g);
                }
                sm.start[b] = v;
                run += c;
            }
        }
        if (tid == 0) {
            const int over = static_cast<int>(ne) + g_all - static_cast<int>(C);   // cells past the last bucket: they wrap to the front
            wrap_s = over > 0 ? static_cast<uint32_t>(over) : 0u;
        }
        __syncthreads();
        mark(3);
        // E4: the cluster that wraps swallows the front buckets it reaches: everything before the first bucket o with g_o > w
        if (tid == 0) {
            uint32_t o = 0;
            if (wrap_s) {
                for (; o < C; o++) {
                    const uint32_t s0 = sm.start[o] & kVal, s1 = o + 1 < C ? (sm.start[o + 1] & kVal) : ne;
                    if (s1 == s0) continue;
                    if (static_cast<int>(o) - static_cast<int>(s0) > static_cast<int>(wrap_s)) break;
                    sm.start[o] = static_cast<uint16_t>(s0);   // no longer a head
                }
            }
        }
        __syncthreads();
        mark(4);
        // E5: placement (any order inside a bucket); start[b] ends up as the END of bucket b, the head bit stays
        for (uint32_t s = tid; s < ne; s += nthreads) {
            const uint32_t b = sm.stash[s];
            const uint32_t old = atomicAdd(packed + (b >> 1), (b & 1u) ? 65536u : 1u);
            const uint32_t t = ((b & 1u) ? (old >> 16) : old) & kVal;
            sm.seq[t] = static_cast<uint16_t>(s);
        }
        __syncthreads();
        const uint32_t w = wrap_s;
        mark(5);
        // A: label every sorted position with its cluster = the sorted position of the cluster's first element (the first
        // element of a head bucket): a prefix maximum over contiguous chunks. Positions before the first head belong to the
        // cluster that wraps = the last head.
        {
            const uint32_t q = robin_block_chunk(ne, nthreads);
            const uint32_t t_lo = min(ne, tid * q), t_hi = min(ne, t_lo + q);
            unsigned long long firsts = 0;
            int last_first = INT_MIN;
            for (uint32_t t = t_lo; t < t_hi; t++) {
                const uint32_t b = sm.stash[sm.seq[t]];
                if ((sm.start[b] & kHead) && t == (b ? static_cast<uint32_t>(sm.start[b - 1] & kVal) : 0u)) {
                    firsts |= 1ull << (t - t_lo);
                    last_first = static_cast<int>(t);
                }
            }
            uint32_t unused_b, unused_a;
            int before, overall;
            robin_block_scan(0u, last_first, unused_b, before, unused_a, overall, 0);
            int label = before == INT_MIN ? overall : before;
            for (uint32_t t = t_lo; t < t_hi; t++) {
                if ((firsts >> (t - t_lo)) & 1ull) label = static_cast<int>(t);
                sm.cells[t] = static_cast<uint16_t>(label);
            }
        }
        __syncthreads();
        mark(6);
        // `start` is dead from here: [0, n) = ideal cell of the cluster's k-th insertion, [n, 2n) = the empty cell it filled,
        // both at sorted position (t0 + k) mod ne (cell i of a cluster = table position (b0 + i) mod C = sorted position
        // (t0 + i) mod ne, so the cluster's cells and its insertions share one index range)
        uint16_t* const atime = sm.start;
        uint16_t* const etime = sm.start + kRobinBlockMax;
        // R2: cluster lengths: every cluster's first element closes the cluster before it; the length replaces the (redundant)
        // label of that cluster's first element: cells[t0] = kHead | length. Readers accept either form of a first element.
        for (uint32_t t = tid; t < ne; t += nthreads) {
            const uint32_t v = sm.cells[t];
            if (!(v & kHead) && v != t) continue;
            const uint32_t pv = sm.cells[t ? t - 1 : ne - 1], t_last = t ? t - 1 : ne - 1;
            const uint32_t t_prev = (pv & kHead) ? t_last : pv;
            const uint32_t len = t_prev == t ? ne : (t > t_prev ? t - t_prev : t + ne - t_prev);
            sm.cells[t_prev] = static_cast<uint16_t>(kHead | len);
            if (len > kRobinBlockMaxCluster) refused_s = 1;
        }
        __syncthreads();
        if (refused_s) return false;
        mark(7);
        auto cluster_of = [&](uint32_t t, uint32_t& t0, uint32_t& L) {
            const uint32_t v = sm.cells[t];
            t0 = (v & kHead) ? t : v;
            L = (v & kHead) ? (v & kVal) : (sm.cells[t0] & kVal);
        };
        // R3: rank of every member among its cluster's members by sequence index = its place in the cluster's insertion
        // history; its ideal cell goes to that place
        {
            int k = 0;
            for (uint32_t t = tid; t < ne; t += nthreads, k++) {
                uint32_t t0, L;
                cluster_of(t, t0, L);
                const uint32_t mine = sm.seq[t];
                const uint32_t a = (static_cast<uint32_t>(sm.stash[mine]) - sm.stash[sm.seq[t0]]) & mask;
                uint32_t r = 0;
#pragma unroll 1   // lanes of a warp walk clusters of different lengths: a tight loop, not unrolled variants of it
                for (uint32_t i = 0; i < L; i++) r += sm.seq[at(t0 + i)] < mine ? 1u : 0u;
                l_r[k] = static_cast<uint16_t>(r);
                atime[at(t0 + r)] = static_cast<uint16_t>(a);
            }
        }
        __syncthreads();
        mark(8);
        // P: the empty cell every insertion filled = the first empty cell at or after its ideal cell (robin_hash.h:1226-1317
        // probes on from there, whatever it swaps on the way). Serial per cluster, by the thread of its first element.
        for (uint32_t t0 = tid; t0 < ne; t0 += nthreads) {
            const uint32_t v = sm.cells[t0];
            if (!(v & kHead)) continue;
            const uint32_t L = v & kVal;
            if (L <= 64) {
                unsigned long long occupied = 0;
#pragma unroll 1
                for (uint32_t k = 0; k < L; k++) {
                    const uint32_t u = at(t0 + k), a = atime[u];
                    const uint32_t e = a + static_cast<uint32_t>(__ffsll(static_cast<long long>(~occupied >> a))) - 1u;
                    occupied |= 1ull << e;
                    etime[u] = static_cast<uint16_t>(e);
                }
            } else {   // a long cluster: occupancy of cell i in bit 15 of atime[t0 + i]
#pragma unroll 1
                for (uint32_t k = 0; k < L; k++) {
                    const uint32_t u = at(t0 + k);
                    uint32_t e = atime[u] & kVal;
                    while (atime[at(t0 + e)] & kHead) e++;
                    atime[at(t0 + e)] |= kHead;
                    etime[u] = static_cast<uint16_t>(e);
                }
            }
        }
        __syncthreads();
        mark(9);
        // W: every member follows its own group through the cluster's history. Before its own insertion it counts the
        // earlier arrivals of its group (it is appended behind them); after it, a later arrival of the group makes the group
        // longer, and an insertion (a2, e2) with a2 < a < e2 carries the group's first element to its end. The group's first
        // cell is the number of members with a smaller ideal cell. The result goes to `stash` (dead since R3).
        {
            int k = 0;
            for (uint32_t t = tid; t < ne; t += nthreads, k++) {
                uint32_t t0, L;
                cluster_of(t, t0, L);
                const uint32_t r = l_r[k], a = atime[at(t0 + r)] & kVal;
                uint32_t before = 0, idx = 0;
#pragma unroll 1
                for (uint32_t i = 0; i < r; i++) {
                    const uint32_t a2 = atime[at(t0 + i)] & kVal;
                    before += a2 < a ? 1u : 0u;
                    idx += a2 == a ? 1u : 0u;
                }
                uint32_t size = idx + 1u;
#pragma unroll 1
                for (uint32_t i = r + 1u; i < L; i++) {
                    const uint32_t u = at(t0 + i), a2 = atime[u] & kVal, e2 = etime[u];
                    before += a2 < a ? 1u : 0u;
                    if (a2 == a) size++;
                    else if (a2 < a && a < e2) idx = idx ? idx - 1u : size - 1u;
                }
                sm.stash[at(t0 + before + idx)] = static_cast<uint16_t>(sm.seq[t]);
            }
        }
        __syncthreads();
        mark(10);
        // E7: the table's iteration order: cell order, the w wrapped cells first
        const bool last = hi == n;
        for (uint32_t t = tid; t < ne; t += nthreads) {
            uint32_t d = t + w;
            if (d >= ne) d -= ne;
            const uint32_t s = sm.stash[t];
            const uint32_t e = s < lo ? sm.prev[s] : s;
            if (last) out[d] = ids ? __ldg(ids + e) : e;
            else sm.seq[d] = static_cast<uint16_t>(e);
        }
        __syncthreads();
        if (!last) {
            for (uint32_t d = tid; d < ne; d += nthreads) sm.prev[d] = sm.seq[d];   // ne <= n / 2 whenever another epoch follows
            __syncthreads();
        }
        mark(13);
        if (prof && threadIdx.x == 0) {
            long long all = 0;
            for (int k = 1; k <= 13; k++) all += prof[k];
            prof[32 + j] = all - prof[31];
            prof[31] = all;
        }
        lo = hi;
    }
    return true;
}

}  // namespace lo
