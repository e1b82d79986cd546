// rtb_sort.cuh -- device LSD radix sort of (Morton key, primitive index) pairs, 8 bits per pass.
//
// Hand-written (no CUB/Thrust).  Per pass: (1) per-tile digit histogram, (2) exclusive scan of the
// digit-major histogram table, (3) stable scatter.  The scatter ranks keys inside a tile with
// warp-synchronous ballot multi-split: each warp owns 32*KPT consecutive keys, ranks them
// round by round against per-warp digit counters in shared memory, then the 256 digit columns are
// prefix-summed across the tile's warps; the tile is then put in digit order in shared memory and written out as runs.
// HBM-bound streaming work: 2 key reads + 1 pair read/write per pass.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

constexpr int RS_TILE = 4096;                  // keys per block (histogram and scatter use the same tiles)
constexpr int RS_HTHREADS = 256;               // histogram: 256 threads x 16 keys
constexpr int RS_HKPT = RS_TILE / RS_HTHREADS;
constexpr int RS_THREADS = 512;                // scatter: 512 threads x 8 keys -- 16 warps per CTA, 8 ranking rounds per warp
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_KPT = RS_TILE / RS_THREADS;

template <typename K>
__global__ void __launch_bounds__(RS_HTHREADS) rs_hist(const K* __restrict__ keys, int n, int shift, int nblocks,
                                                       unsigned* __restrict__ hist)
{
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * RS_TILE;
    K k[RS_HKPT];
#pragma unroll
    for (int r = 0; r < RS_HKPT; r++) {            // all loads in flight before the first shared-memory atomic
        int i = base + r * RS_HTHREADS + threadIdx.x;
        k[r] = i < n ? keys[i] : (K)0;
    }
#pragma unroll
    for (int r = 0; r < RS_HKPT; r++) {
        int i = base + r * RS_HTHREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(k[r] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Exclusive scan of the digit-major histogram table, in two levels so that it runs on 256 CTAs instead of one:
// rs_scan_rows: CTA d scans row d (the tile counts of digit d) in place and writes the row total;
// the 256 row totals are prefix-summed by every scatter CTA itself (one warp-shuffle scan, rs_scatter).
__global__ void __launch_bounds__(256) rs_scan_rows(unsigned* __restrict__ data, int nblocks, unsigned* __restrict__ totals)
{
    __shared__ unsigned part[256];
    unsigned* row = data + (size_t)blockIdx.x * nblocks;
    const int t = threadIdx.x;
    const int per = (nblocks + 255) / 256;
    const int b = min(t * per, nblocks), e = min(b + per, nblocks);
    unsigned s = 0;
    for (int i = b; i < e; i++) s += row[i];
    part[t] = s;
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {          // Hillis-Steele inclusive scan
        unsigned v = t >= off ? part[t - off] : 0u;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    unsigned run = part[t] - s;
    for (int i = b; i < e; i++) { unsigned v = row[i]; row[i] = run; run += v; }
    if (t == 255) totals[blockIdx.x] = part[255];
}

// The tile is first sorted by digit INSIDE shared memory (rank -> local position), then written out in local order, so
// consecutive threads store consecutive addresses of one digit's run (16 keys on average) instead of 32 lanes hitting 32
// different bins with 8-byte stores (the first version: 1.2 TB/s of DRAM traffic, 195 us per pass on 10 M 64-bit keys).
template <typename K>
__global__ void __launch_bounds__(RS_THREADS, 2) rs_scatter(const K* __restrict__ keys_in, const unsigned* __restrict__ vals_in,
                                                         K* __restrict__ keys_out, unsigned* __restrict__ vals_out, int n,
                                                         int shift, int nblocks, const unsigned* __restrict__ offs,
                                                         const unsigned* __restrict__ totals)
{
    __shared__ unsigned short cnt[RS_WARPS][256];   // per warp and digit: count, then first local position (<= 4096)
    __shared__ unsigned dbase[256];                 // exclusive prefix of the 256 digit totals, then (global base - local base) per digit
    __shared__ K buf[RS_TILE];                      // the tile in digit order: keys, then (re-used) values
    __shared__ unsigned char dig[RS_TILE];          // digit of the element at each local position
    __shared__ unsigned wsum[8];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool digit_thread = threadIdx.x < 256;    // threads 0..255 (warps 0..7) each own one digit in the prefix steps
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&cnt[0][0])[i] = 0;
    {   // thread d: exclusive prefix of totals[0..d): warp scan + warp sums
        const unsigned v = digit_thread ? totals[threadIdx.x] : 0u;
        unsigned inc = v;
        for (int o = 1; o < 32; o <<= 1) { unsigned u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (digit_thread && lane == 31) wsum[w] = inc;
        __syncthreads();
        if (digit_thread) {
            unsigned ws = 0;
            for (int ww = 0; ww < w; ww++) ws += wsum[ww];
            dbase[threadIdx.x] = ws + inc - v;
        }
    }
    __syncthreads();
    const int tbase = blockIdx.x * RS_TILE;
    const int tile_n = min(RS_TILE, n - tbase);
    const int wbase = tbase + w * (32 * RS_KPT);
    K key[RS_KPT];
    unsigned val[RS_KPT];                           // loaded with the keys: one exposed DRAM latency per tile, not two
    unsigned short off[RS_KPT];
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RS_KPT; r++) {
        int i = wbase + r * 32 + lane;
        key[r] = i < n ? keys_in[i] : (K)0;
    }
#pragma unroll
    for (int r = 0; r < RS_KPT; r++) {
        int i = wbase + r * 32 + lane;
        val[r] = i < n ? vals_in[i] : 0u;
    }
#pragma unroll
    for (int r = 0; r < RS_KPT; r++) {
        int i = wbase + r * 32 + lane;
        unsigned d = i < n ? ((unsigned)(key[r] >> shift) & 255u) : 256u;
        // lanes holding the same digit: eight ballots instead of MATCH.ANY, whose issue rate (not its latency) bounded this
        // kernel -- 2 M MATCH instructions per pass took longer than moving the data (profiles/r01f_ab_experiments.txt)
        unsigned peers = __ballot_sync(0xffffffffu, d < 256u);
        if (d >= 256u) peers = ~peers;
#pragma unroll
        for (int b = 0; b < 8; b++) {
            const unsigned m = __ballot_sync(0xffffffffu, (d >> b) & 1u);
            peers &= ((d >> b) & 1u) ? m : ~m;
        }
        int leader = __ffs(peers) - 1;
        unsigned old = 0;
        if (lane == leader && d < 256u) { old = cnt[w][d]; cnt[w][d] = (unsigned short)(old + __popc(peers)); }
        old = __shfl_sync(0xffffffffu, old, leader);
        off[r] = (unsigned short)(old + __popc(peers & lt));
        __syncwarp();
    }
    __syncthreads();
    {   // digit d (= thread d): its count in the tile, the tile-local exclusive prefix over digits, the per-warp local bases
        const int d = threadIdx.x & 255;
        unsigned tot = 0;
        if (digit_thread) {
#pragma unroll
            for (int ww = 0; ww < RS_WARPS; ww++) tot += cnt[ww][d];
        }
        unsigned inc = tot;
        for (int o = 1; o < 32; o <<= 1) { unsigned u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        __syncthreads();                            // wsum is free again (every digit thread read it above)
        if (digit_thread && lane == 31) wsum[w] = inc;
        __syncthreads();
        if (digit_thread) {
            unsigned ws = 0;
            for (int ww = 0; ww < w; ww++) ws += wsum[ww];
            const unsigned local = ws + inc - tot;  // first local position of digit d
            unsigned run = local;
#pragma unroll
            for (int ww = 0; ww < RS_WARPS; ww++) { unsigned c = cnt[ww][d]; cnt[ww][d] = (unsigned short)run; run += c; }
            dbase[d] = offs[(size_t)d * nblocks + blockIdx.x] + dbase[d] - local;  // global position = dbase[d] + local position
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_KPT; r++) {
        int i = wbase + r * 32 + lane;
        if (i < n) {
            unsigned d = (unsigned)(key[r] >> shift) & 255u;
            unsigned pos = (unsigned)cnt[w][d] + off[r];
            off[r] = (unsigned short)pos;
            buf[pos] = key[r];
            dig[pos] = (unsigned char)d;
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < tile_n; j += RS_THREADS) keys_out[dbase[dig[j]] + j] = buf[j];
    __syncthreads();
    unsigned* vbuf = (unsigned*)buf;
#pragma unroll
    for (int r = 0; r < RS_KPT; r++) {
        int i = wbase + r * 32 + lane;
        if (i < n) vbuf[off[r]] = val[r];
    }
    __syncthreads();
    for (int j = threadIdx.x; j < tile_n; j += RS_THREADS) vals_out[dbase[dig[j]] + j] = vbuf[j];
}

// `hist` holds 256 * nblocks + 256 counters.  Sorts `bits` low bits; result ends in (keys[0], vals[0]) or (keys[1], vals[1]); returns which.
template <typename K>
inline int radix_sort_pairs(K* keys[2], unsigned* vals[2], int n, int bits, unsigned* hist, cudaStream_t st)
{
    const int nblocks = (n + RS_TILE - 1) / RS_TILE;
    int cur = 0;
    for (int shift = 0; shift < bits; shift += 8) {
        rs_hist<K><<<nblocks, RS_HTHREADS, 0, st>>>(keys[cur], n, shift, nblocks, hist);
        rs_scan_rows<<<256, 256, 0, st>>>(hist, nblocks, hist + (size_t)256 * nblocks);
        rs_scatter<K><<<nblocks, RS_THREADS, 0, st>>>(keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, shift, nblocks, hist,
                                                      hist + (size_t)256 * nblocks);
        cur ^= 1;
    }
    return cur;
}

}  // namespace rtb
