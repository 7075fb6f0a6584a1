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
// rises), then one thread per cluster replays the literal robin-hood insertion of its members, in sequence order, on the
// cluster's handful of cells. The iteration order of the epoch's table (the clusters' cells in position order; the w cells
// that ran past the last bucket come first) is the next epoch's re-insertion sequence, and the last one is the answer.
// Identical to the sequential replay of lo_robin.cu, which tests/test_gpu_robin.py pins to the real container.
#pragma once

#include <limits.h>
#include <stdint.h>

namespace lo {

constexpr uint32_t kRobinBlockMax = 16384;          // elements the shared-memory replay takes (a VLP16 update set is 12-15 k)
constexpr uint32_t kRobinBlockBuckets = 2 * kRobinBlockMax;
// A cluster longer than this is refused (the function returns false): one thread replays a cluster, so a long one would
// crawl, and past 8192 cells tsl's probe-distance limit (robin_hash.h:327) adds growth steps this replay does not model.
// Hash-random keys at load <= 0.5 give clusters of tens of cells; the sequential replay of lo_robin.cu takes anything.
constexpr uint32_t kRobinBlockMaxCluster = 2048;
// uint16 arrays: start[C] | seq[n] | stash[n] | cells[n] | prev[n / 2]
constexpr size_t kRobinBlockSmemBytes = 2 * (kRobinBlockBuckets + 3 * kRobinBlockMax + kRobinBlockMax / 2);

struct RobinBlockSmem {
    uint16_t* start;   // [C]  counts -> exclusive starts (bit 15: the bucket heads a cluster) -> ends after the placement
    uint16_t* seq;     // [n]  sequence indices grouped by bucket, then each cluster's in sequence order; E7: scratch
    uint16_t* stash;   // [n]  ideal bucket of sequence element s
    uint16_t* cells;   // [n]  the occupied cells in position order from the first non-wrapped one: sequence index of the resident
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

// exclusive prefix sum and exclusive prefix maximum of one value per thread, and the block totals; blockDim.x <= 1024
__device__ inline void robin_block_scan(uint32_t v, int g, uint32_t& v_before, int& g_before, uint32_t& v_all, int& g_all) {
    __shared__ uint32_t ws[32];
    __shared__ int wg[32];
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
        ws[warp] = vi;
        wg[warp] = gi;
    }
    __syncthreads();
    uint32_t vb = 0, va = 0;
    int gb = INT_MIN, ga = INT_MIN;
    for (int w = 0; w < nwarps; w++) {
        if (w < warp) {
            vb += ws[w];
            gb = max(gb, wg[w]);
        }
        va += ws[w];
        ga = max(ga, wg[w]);
    }
    const uint32_t v_excl_in_warp = __shfl_up_sync(0xFFFFFFFFu, vi, 1);
    const int g_excl_in_warp = __shfl_up_sync(0xFFFFFFFFu, gi, 1);
    v_before = vb + (lane ? v_excl_in_warp : 0u);
    g_before = lane ? max(gb, g_excl_in_warp) : gb;
    v_all = va;
    g_all = ga;
    __syncthreads();
}

// uh[k]: the reference's 22-bit hash of the k-th key in insertion (arrival) order, k < n <= kRobinBlockMax (global memory).
// ids[k]: what to emit for element k. out[d] = ids[element at iteration position d]. All threads of ONE block call.
// Returns false (in every thread; `out` is then unspecified) if a cluster exceeded kRobinBlockMaxCluster.
__device__ inline bool robin_order_block(const uint32_t* __restrict__ uh, const uint32_t* __restrict__ ids, uint32_t n, unsigned char* smem,
                                         uint32_t* __restrict__ out) {
    const RobinBlockSmem sm = robin_block_carve(smem);
    __shared__ uint32_t wrap_s, origin_s;
    __shared__ int refused_s;
    if (threadIdx.x == 0) refused_s = 0;
    __syncthreads();
    const uint32_t tid = threadIdx.x, nthreads = blockDim.x;
    constexpr uint16_t kEmpty = 0xFFFFu, kHead = 0x8000u, kVal = 0x7FFFu;
    uint32_t lo = 0;   // elements present before this epoch's arrivals = length of `prev`
    for (uint32_t j = 0; lo < n; j++) {
        const uint32_t C = 2u << j, mask = C - 1;
        const uint32_t hi = j == 0 ? 1u : min(n, 1u << j);   // arrivals of this epoch: element ids lo .. hi - 1
        const uint32_t ne = hi;                               // sequence S = prev[0 .. lo) ++ [lo .. hi): S[s] = s for s >= lo
        uint32_t* packed = reinterpret_cast<uint32_t*>(sm.start);
        // E1: clear the counts
        for (uint32_t w = tid; w < (C + 1) / 2; w += nthreads) packed[w] = 0;
        __syncthreads();
        // E2: histogram of the ideal buckets (two 16-bit counters per word; a count is at most ne < 2^15)
        for (uint32_t s = tid; s < ne; s += nthreads) {
            const uint32_t e = s < lo ? sm.prev[s] : s;
            const uint32_t b = __ldg(uh + e) & mask;
            sm.stash[s] = static_cast<uint16_t>(b);
            atomicAdd(packed + (b >> 1), (b & 1u) ? 65536u : 1u);
        }
        __syncthreads();
        // E3: counts -> exclusive starts, cluster heads, overflow past the last bucket. Thread t owns buckets [t per, (t+1) per).
        const uint32_t per = (C + nthreads - 1) / nthreads;
        const uint32_t b_lo = min(C, tid * per), b_hi = min(C, b_lo + per);
        uint32_t sum = 0;
        for (uint32_t b = b_lo; b < b_hi; b++) sum += sm.start[b];
        uint32_t base, unused_all;
        int unused_g, unused_ga;
        robin_block_scan(sum, INT_MIN, base, unused_g, unused_all, unused_ga);
        int gmax = INT_MIN;   // g_b = b - start_b over my non-empty buckets
        {
            uint32_t run = base;
            for (uint32_t b = b_lo; b < b_hi; b++) {
                const uint32_t c = sm.start[b];
                if (c) gmax = max(gmax, static_cast<int>(b) - static_cast<int>(run));
                run += c;
            }
        }
        uint32_t unused_vb, unused_va;
        int g_before, g_all;
        robin_block_scan(0u, gmax, unused_vb, g_before, unused_va, g_all);
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
            origin_s = o;
        }
        __syncthreads();
        // E5: placement (any order inside a bucket); start[b] ends up as the END of bucket b, the head bit stays
        for (uint32_t s = tid; s < ne; s += nthreads) {
            const uint32_t b = sm.stash[s];
            const uint32_t old = atomicAdd(packed + (b >> 1), (b & 1u) ? 65536u : 1u);
            const uint32_t t = ((b & 1u) ? (old >> 16) : old) & kVal;
            sm.seq[t] = static_cast<uint16_t>(s);
        }
        __syncthreads();
        const uint32_t w = wrap_s, origin = origin_s;
        auto bucket_begin = [&](uint32_t b) { return b ? static_cast<uint32_t>(sm.start[b - 1] & kVal) : 0u; };
        auto bucket_end = [&](uint32_t b) { return static_cast<uint32_t>(sm.start[b] & kVal); };
        // E6a: every cluster head labels its members with the cluster's first sorted index (in `cells`, free until E6c)
        for (uint32_t b0 = b_lo; b0 < b_hi; b0++) {
            if (!(sm.start[b0] & kHead)) continue;
            const uint32_t t0 = bucket_begin(b0);
            uint32_t b = b0;
            bool wrapped = false;
            while (true) {
                for (uint32_t t = bucket_begin(b); t < bucket_end(b); t++) sm.cells[t] = static_cast<uint16_t>(t0);
                b++;
                if (b == C) {
                    if (!w) break;     // nothing ran past the last bucket
                    b = 0;             // the wrapping cluster goes on through the front buckets up to `origin`
                    wrapped = true;
                }
                if (wrapped ? (b >= origin || b == b0) : (sm.start[b] & kHead) != 0) break;
            }
        }
        __syncthreads();
        // E6b: every member finds its rank by sequence index inside its cluster (all members in parallel; clusters are short)
        constexpr int kMaxOwn = (kRobinBlockMax + 255) / 256;   // members per thread at >= 256 threads
        uint16_t own_rank[kMaxOwn];
        {
            int k = 0;
            for (uint32_t t = tid; t < ne; t += nthreads, k++) {
                const uint32_t t0 = sm.cells[t], mine = sm.seq[t];
                uint32_t r = 0, u = t0;
                for (uint32_t steps = 0; steps < ne; steps++) {
                    if (sm.cells[u] != t0) break;
                    r += sm.seq[u] < mine ? 1u : 0u;
                    u = u + 1 == ne ? 0u : u + 1;
                }
                own_rank[k] = static_cast<uint16_t>(r);
            }
        }
        __syncthreads();
        {
            int k = 0;
            uint16_t own_seq[kMaxOwn];
            for (uint32_t t = tid; t < ne; t += nthreads, k++) own_seq[k] = sm.seq[t];
            __syncthreads();
            k = 0;
            for (uint32_t t = tid; t < ne; t += nthreads, k++) {
                uint32_t at = static_cast<uint32_t>(sm.cells[t]) + own_rank[k];
                if (at >= ne) at -= ne;
                sm.seq[at] = own_seq[k];   // a permutation inside the cluster: its members in sequence (= insertion) order
            }
        }
        __syncthreads();
        // E6c: one thread per cluster replays robin_hash's insertion (robin_hash.h:1226-1317) on the cluster's own cells.
        // Cell i of the cluster is table position (b0 + i) mod C and lives at cells[(t0 + i) mod ne]. The lengths are taken
        // first, while the labels are intact everywhere (a neighbour's cells may hold anything once it has started).
        constexpr int kMaxHeads = (kRobinBlockBuckets + 255) / 256;   // buckets per thread at >= 256 threads
        uint16_t own_len[kMaxHeads];
        {
            int k = 0;
            for (uint32_t b0 = b_lo; b0 < b_hi; b0++, k++) {
                own_len[k] = 0;
                if (!(sm.start[b0] & kHead)) continue;
                const uint32_t t0 = bucket_begin(b0);
                uint32_t L = 0;
                for (uint32_t u = t0; L < ne && sm.cells[u] == t0; L++) u = u + 1 == ne ? 0u : u + 1;
                own_len[k] = static_cast<uint16_t>(L);
                if (L > kRobinBlockMaxCluster) refused_s = 1;
            }
        }
        __syncthreads();
        if (refused_s) return false;
        {
            int k = 0;
            for (uint32_t b0 = b_lo; b0 < b_hi; b0++, k++) {
                const uint32_t L = own_len[k];
                if (!L) continue;
                const uint32_t t0 = bucket_begin(b0);
                auto at = [&](uint32_t i) { const uint32_t u = t0 + i; return u >= ne ? u - ne : u; };
                if (L == 1) {
                    sm.cells[t0] = sm.seq[t0];
                    continue;
                }
                for (uint32_t i = 0; i < L; i++) sm.cells[at(i)] = kEmpty;
                for (uint32_t m = 0; m < L; m++) {
                    uint32_t carry = sm.seq[at(m)];
                    uint32_t i = (static_cast<uint32_t>(sm.stash[carry]) - b0) & mask;   // ideal cell, inside the cluster by construction
                    uint32_t dist = 0;
                    while (i < L) {
                        const uint32_t cur = sm.cells[at(i)];
                        if (cur == kEmpty) {
                            sm.cells[at(i)] = static_cast<uint16_t>(carry);
                            break;
                        }
                        const uint32_t cur_dist = (i - ((static_cast<uint32_t>(sm.stash[cur]) - b0) & mask)) & mask;
                        if (dist > cur_dist) {   // a strictly richer resident gives way and is carried on
                            sm.cells[at(i)] = static_cast<uint16_t>(carry);
                            carry = cur;
                            dist = cur_dist;
                        }
                        i++;
                        dist++;
                    }
                }
            }
        }
        __syncthreads();
        // E7: the table's iteration order: cell order, the w wrapped cells first
        const bool last = hi == n;
        for (uint32_t t = tid; t < ne; t += nthreads) {
            uint32_t d = t + w;
            if (d >= ne) d -= ne;
            const uint32_t s = sm.cells[t];
            const uint32_t e = s < lo ? sm.prev[s] : s;
            if (last) out[d] = __ldg(ids + e);
            else sm.seq[d] = static_cast<uint16_t>(e);
        }
        __syncthreads();
        if (!last) {
            for (uint32_t d = tid; d < ne; d += nthreads) sm.prev[d] = sm.seq[d];   // ne <= n / 2 whenever another epoch follows
            __syncthreads();
        }
        lo = hi;
    }
    return true;
}

}  // namespace lo
