// sse_kernels.cuh -- SSE event scan over the per-stream bodies the demux gathered.
// The tunnel forwards server-sent events untouched ("sse" feature, /root/reference/tunnel/src/protocol.rs:36,68);
// the framing is the upstream's: `data: {...}\n\n` ... `data: [DONE]\n\n` (/root/reference/tmp/mock_llm.py:54,68-69).
// One warp per stream walks the body in 512-byte rows (aligned 16-byte loads, bytes outside the stream read
// as zero), marks every position that closes a "\n\n" pair, and either counts them (pass 1) or writes the
// event end offsets in order (pass 2).  HBM-bound: every body byte is read once per pass.
#pragma once
#include "frame_kernels.cuh"

namespace p2pt {

// 0x80 in every byte b of the result whose byte in w is '\n' and whose predecessor (byte b - 1, the top byte of
// prev_word for b = 0) is '\n' too - the byte that closes an event.  One funnel shift lines the predecessors up, one
// LOP3 ORs the two differences from 0x0a, and an exact SWAR zero-byte test (no borrow between bytes) flags the bytes
// where both vanish: five instructions per word (testing for '\n' first and pairing the flags afterwards took eight,
// and the scan kernels were issue-bound at 68 / 135 instructions per 512-byte row).
__device__ __forceinline__ uint32_t lf_pair_flags(uint32_t w, uint32_t prev_word) {
    const uint32_t s = __funnelshift_l(prev_word, w, 8);                // bytes [prev.3, w.0, w.1, w.2]
    const uint32_t z = (w ^ 0x0a0a0a0au) | (s ^ 0x0a0a0a0au);           // zero byte <=> both are '\n'
    return ~(((z & 0x7f7f7f7fu) + 0x7f7f7f7fu) | z) & 0x80808080u;
}
struct PairFlags { uint32_t f[4]; };
__device__ __forceinline__ PairFlags lf_pairs16(const uint4& v, uint32_t prev_word) {
    PairFlags r;
    r.f[0] = lf_pair_flags(v.x, prev_word); r.f[1] = lf_pair_flags(v.y, v.x);
    r.f[2] = lf_pair_flags(v.z, v.y);       r.f[3] = lf_pair_flags(v.w, v.z);
    return r;
}
__device__ __forceinline__ uint32_t pair_count(const PairFlags& p) { return __popc(p.f[0]) + __popc(p.f[1]) + __popc(p.f[2]) + __popc(p.f[3]); }
// bit b = byte b of the vector closes an event (only lanes that hold an event pay for this: the multiply gathers the
// four flag bits of a word into one nibble)
__device__ __forceinline__ uint32_t pair_mask16(const PairFlags& p) {
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) m |= (((p.f[k] >> 7) * 0x01020408u) >> 24) << (4 * k);
    return m;
}

// keep bytes [lo, hi) of v, zero the others (0 <= lo <= hi <= 16); zero is not '\n'
__device__ __forceinline__ uint4 keep_bytes16(const uint4& v, int lo, int hi) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int l = lo - 4 * k, h = hi - 4 * k;                        // range inside this word
        const uint32_t ml = l <= 0 ? 0xffffffffu : l >= 4 ? 0u : (0xffffffffu << (8 * l));
        const uint32_t mh = h >= 4 ? 0xffffffffu : h <= 0 ? 0u : (0xffffffffu >> (32 - 8 * h));
        w[k] &= ml & mh;
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// does an event that starts at body + pos read `data: [DONE]\n\n`?  Byte 6 tells `data: [` from `data: {`,
// so ordinary events cost one byte load.
__device__ __forceinline__ bool sse_is_done_event(const uint8_t* __restrict__ body, u64 pos, u64 stream_end) {
    if (stream_end - pos < 14 || __ldg(body + pos + 6) != (uint8_t)'[') return false;
    const char* D = "data: [DONE]\n\n";
    bool eq = true;
    for (int q = 0; q < 14; q++) eq &= (__ldg(body + pos + q) == (uint8_t)D[q]);
    return eq;
}

// (Splitting every stream into four parts, one warp each, to raise occupancy on a few thousand long streams
// measured the same 0.78-0.79 ms per 1.2 GB: the passes are issue-bound on the per-byte tests, not short of warps.)
// One warp per stream; the stream is walked in groups of kSseRows rows of 512 bytes whose ALIGNED 16-byte
// loads are all issued before the first is used (one vector per lane and row; a vector that sticks out of
// the stream is masked, never byte-loaded).  The count pass only sums; the emit pass needs each lane's
// exclusive event index, two rows share one 5-step shuffle scan (16-bit fields).
#ifndef P2PT_SSE_ROWS
#define P2PT_SSE_ROWS 4          // measured per 1.2 GB of bodies: 2 rows 0.84 ms, 4 rows 0.79 ms, 8 rows 0.91 ms (80 registers)
#endif
constexpr int kSseRows = P2PT_SSE_ROWS;

template <bool EMIT>
__global__ void __launch_bounds__(256) sse_scan_kernel(u64 k, const uint8_t* __restrict__ body, const u64* __restrict__ body_off,
                                                      u64* __restrict__ ev_csr, u64* __restrict__ ev_end, u64 cap,
                                                      uint8_t* __restrict__ done) {
    const int lane = threadIdx.x & 31;
    const u64 warp = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((u64)gridDim.x * blockDim.x) >> 5;
    for (u64 s = warp; s < k; s += nwarps) {
        const u64 a = __ldg(body_off + s), b = __ldg(body_off + s + 1);
        const uint8_t* lo = body + a; const uint8_t* hi = body + b;
        const uint8_t* row0 = lo - ((uintptr_t)lo & 15);                  // aligned start of the first row
        u64 count = 0;
        uint32_t lane_count = 0;
        const u64 out0 = EMIT ? ev_csr[s] : 0;
        uint32_t carry = 0;                                   // last word of the previous row (0 at the stream start)
        bool is_done = false;
        if (EMIT && lane == 0) is_done = sse_is_done_event(body, a, b);   // an event starts at the stream start
        for (const uint8_t* rp = row0; rp < hi; rp += 512 * kSseRows) {
            uint4 v[kSseRows];
#pragma unroll
            for (int r = 0; r < kSseRows; r++) {
                const uint8_t* p = rp + 512 * r + 16 * lane;
                v[r] = (p < hi) ? ldg128(p) : make_uint4(0, 0, 0, 0);
            }
            PairFlags m[kSseRows];
            uint32_t c[kSseRows];
#pragma unroll
            for (int r = 0; r < kSseRows; r++) {
                const uint8_t* p = rp + 512 * r + 16 * lane;
                if (p < hi && (p < lo || p + 16 > hi))
                    v[r] = keep_bytes16(v[r], p < lo ? (int)(lo - p) : 0, p + 16 > hi ? (int)(hi - p) : 16);
                uint32_t prev = __shfl_up_sync(0xffffffffu, v[r].w, 1);
                if (lane == 0) prev = carry;
                carry = __shfl_sync(0xffffffffu, v[r].w, 31);
                m[r] = lf_pairs16(v[r], prev);
                c[r] = pair_count(m[r]);
            }
            if (!EMIT) {
#pragma unroll
                for (int r = 0; r < kSseRows; r++) lane_count += c[r];
            } else if (!__any_sync(0xffffffffu, (c[0] | c[1] | c[2] | c[3]) > 1u)) {
                // the usual case - no lane holds two events of a row (events are longer than 16 bytes): a lane's event
                // index is a ballot and a popcount, no shuffle scan, and there is no loop over the lane's events
                static_assert(kSseRows == 4, "four rows per group");
#pragma unroll
                for (int r = 0; r < kSseRows; r++) {
                    const uint32_t bal = __ballot_sync(0xffffffffu, c[r] != 0);
                    if (c[r]) {
                        const u64 slot = out0 + count + __popc(bal & ((1u << lane) - 1u));
                        const u64 end = (u64)(rp + 512 * r + 16 * lane - body) + __ffs(pair_mask16(m[r]));   // just past the second '\n'
                        if (slot < cap) ev_end[slot] = end;
                        is_done |= sse_is_done_event(body, end, b);            // does the next event read `data: [DONE]`?
                    }
                    count += __popc(bal);
                }
            } else {
#pragma unroll
                for (int r2 = 0; r2 < kSseRows; r2 += 2) {
                    // a row holds at most 512 events, so two rows' inclusive scans fit 16-bit fields
                    const uint32_t both = c[r2] | (c[r2 + 1] << 16);
                    uint32_t incl = both;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
                    const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const int r = r2 + e;
                        const uint32_t inc_r = (incl >> (16 * e)) & 0xffffu;
                        u64 slot = out0 + count + (inc_r - c[r]);
                        uint32_t mm = c[r] ? pair_mask16(m[r]) : 0u;
                        const u64 pos0 = (u64)(rp + 512 * r + 16 * lane - body);
                        while (mm) {
                            const int bit = __ffs(mm) - 1; mm &= mm - 1;
                            const u64 end = pos0 + bit + 1;                    // offset just past the second '\n'
                            if (slot < cap) ev_end[slot] = end;
                            slot++;
                            is_done |= sse_is_done_event(body, end, b);        // does the next event read `data: [DONE]`?
                        }
                        count += (tot >> (16 * e)) & 0xffffu;
                    }
                }
            }
        }
        if (!EMIT) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) lane_count += __shfl_xor_sync(0xffffffffu, lane_count, o);
            if (lane == 0) ev_csr[s] = lane_count;             // per-stream counts, scanned afterwards
        } else {
            const bool any = __any_sync(0xffffffffu, is_done);
            if (lane == 0) done[s] = any ? 1 : 0;
        }
    }
}

// exclusive scan of k u64 counts in place (k + 1 entries out): block sums -> single block scan -> add
__global__ void __launch_bounds__(kScanThreads) u64_scan_blocksum_kernel(u64 k, const u64* __restrict__ v, u64* __restrict__ blk) {
    const u64 i = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes x; x.heads = 0; x.bytes = i < k ? v[i] : 0;
    HeadBytes tot;
    (void)block_scan_excl(x, tot);
    if (threadIdx.x == 0) blk[blockIdx.x] = tot.bytes;
}
__global__ void __launch_bounds__(kScanThreads) u64_scan_blocksums_kernel(u64 nblk, u64* __restrict__ blk) {
    __shared__ u64 carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (u64 b0 = 0; b0 < nblk; b0 += kScanThreads) {
        const u64 b = b0 + threadIdx.x;
        HeadBytes x; x.heads = 0; x.bytes = b < nblk ? blk[b] : 0;
        HeadBytes tot;
        HeadBytes ex = block_scan_excl(x, tot);
        const u64 c = carry;
        if (b < nblk) blk[b] = c + ex.bytes;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + tot.bytes;
        __syncthreads();
    }
    if (threadIdx.x == 0) blk[nblk] = carry;
}
__global__ void __launch_bounds__(kScanThreads) u64_scan_apply_kernel(u64 k, u64 nblk, u64* __restrict__ v, const u64* __restrict__ blk,
                                                                     u64* __restrict__ total_out) {
    const u64 i = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes x; x.heads = 0; x.bytes = i < k ? v[i] : 0;
    HeadBytes tot;
    HeadBytes ex = block_scan_excl(x, tot);
    if (i < k) v[i] = blk[blockIdx.x] + ex.bytes;
    if (i == 0) { v[k] = blk[nblk]; *total_out = blk[nblk]; }
}

}  // namespace p2pt
