// dtls_state_kernels.cuh -- DTLS record-layer bookkeeping in batched form:
//   * datagram -> records          webrtc-dtls 0.10.0 record_layer unpack_datagram (un-vendored)
//   * anti-replay sliding window   replaydetector SlidingWindowDetector, window 64, as driven by
//                                  webrtc-dtls conn.rs handle_incoming_packet: check() before the
//                                  record is opened, accept() only once it authenticated
//   * sequence numbers of outgoing records: every record takes its epoch's counter with fetch_add(1)
//                                  (webrtc-dtls conn.rs, the write path), 48 bits, overflow is an error
// Small integer kernels: one thread per datagram / per session walks its few items exactly like the
// reference does; the parallelism is across datagrams and across sessions.
#pragma once
#include "frame_kernels.cuh"

namespace p2pt {

constexpr u64 kSeq48Max = (1ull << 48) - 1;

// ---- datagram split -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t dgram_walk(const uint8_t* d, u64 len, u64 base, u64* rec_off, uint32_t* rec_dgram,
                                               u64 slot0, u64 cap, uint32_t dgram) {
    u64 off = 0;
    uint32_t k = 0;
    while (off < len) {
        // `<=`: unpack_datagram refuses a trailing bare 13-byte header (a record with no payload) too
        if (len - off <= 13) return 0xffffffffu;                       // ErrInvalidPacketLength: whole datagram dropped
        const u64 plen = 13 + (((u64)__ldg(d + off + 11) << 8) | __ldg(d + off + 12));
        if (off + plen > len) return 0xffffffffu;
        if (rec_off && slot0 + k < cap) { rec_off[slot0 + k] = base + off; rec_dgram[slot0 + k] = dgram; }
        k++;
        off += plen;
    }
    return k;
}

__global__ void __launch_bounds__(kScanThreads) dgram_count_kernel(u64 n, const uint8_t* __restrict__ buf,
                                                                  const u64* __restrict__ dgram_off,
                                                                  uint32_t* __restrict__ count, uint8_t* __restrict__ status,
                                                                  uint32_t* __restrict__ blk) {
    const u64 i = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes v; v.heads = 0; v.bytes = 0;
    if (i < n) {
        const u64 o = __ldg(dgram_off + i), len = __ldg(dgram_off + i + 1) - o;
        const uint32_t k = dgram_walk(buf + o, len, o, nullptr, nullptr, 0, 0, 0);
        const bool bad = (k == 0xffffffffu);
        status[i] = bad ? P2PT_ST_BAD_HEADER : P2PT_ST_OK;
        v.heads = bad ? 0u : k;
        count[i] = v.heads;
    }
    HeadBytes tot;
    (void)block_scan_excl(v, tot);
    if (threadIdx.x == 0) blk[blockIdx.x] = tot.heads;
}

// single block: exclusive scan of u32 block sums in place, total in blk[nblk]
__global__ void __launch_bounds__(kScanThreads) u32_scan_blocksums_kernel(u64 nblk, uint32_t* __restrict__ blk) {
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (u64 b0 = 0; b0 < nblk; b0 += kScanThreads) {
        const u64 b = b0 + threadIdx.x;
        HeadBytes v; v.heads = b < nblk ? blk[b] : 0; v.bytes = 0;
        HeadBytes tot;
        HeadBytes ex = block_scan_excl(v, tot);
        const uint32_t c = carry;
        if (b < nblk) blk[b] = c + ex.heads;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + tot.heads;
        __syncthreads();
    }
    if (threadIdx.x == 0) blk[nblk] = carry;
}

__global__ void __launch_bounds__(kScanThreads) dgram_emit_kernel(u64 n, u64 nblk, const uint8_t* __restrict__ buf,
                                                                 const u64* __restrict__ dgram_off,
                                                                 const uint32_t* __restrict__ count,
                                                                 const uint32_t* __restrict__ blk, u64 cap,
                                                                 u64* __restrict__ rec_off, uint32_t* __restrict__ rec_dgram,
                                                                 uint32_t* __restrict__ n_records_out) {
    const u64 i = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes v; v.heads = i < n ? count[i] : 0; v.bytes = 0;
    HeadBytes tot;
    HeadBytes ex = block_scan_excl(v, tot);
    if (i < n && v.heads) {
        const u64 o = __ldg(dgram_off + i), len = __ldg(dgram_off + i + 1) - o;
        (void)dgram_walk(buf + o, len, o, rec_off, rec_dgram, (u64)blk[blockIdx.x] + ex.heads, cap, (uint32_t)i);
    }
    if (i == 0) {
        const uint32_t total = blk[nblk];
        *n_records_out = total;
        // terminator: end of the last record = end of the last valid datagram's data; records are
        // contiguous inside a datagram, so the caller takes rec_off[j + 1] only within one datagram.
        // For convenience the slot after the last record holds the end of the input.
        if (total <= cap) rec_off[total] = __ldg(dgram_off + n);
    }
}

// ---- anti-replay window -----------------------------------------------------------------------------
// keys for the session grouping: a descriptor array the demux can sort (stream_id = session)
__global__ void __launch_bounds__(256) replay_keys_kernel(u64 n, const uint32_t* __restrict__ session, p2pt_desc* __restrict__ desc) {
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        p2pt_desc d;
        d.pay_off = 0; d.pay_len = 0; d.stream_id = session[i]; d.msg_type = 0; d.status = P2PT_ST_OK;
        for (int q = 0; q < 6; q++) d.pad[q] = 0;
        desc[i] = d;
    }
}

__global__ void __launch_bounds__(256) replay_walk_kernel(const uint32_t* __restrict__ k_ptr, const uint32_t* __restrict__ uniq_session,
                                                         const u64* __restrict__ csr, const uint32_t* __restrict__ perm,
                                                         const u64* __restrict__ seq48, const uint8_t* __restrict__ authentic,
                                                         uint32_t n_sessions, u64* __restrict__ latest, u64* __restrict__ mask,
                                                         uint8_t* __restrict__ started, uint8_t* __restrict__ accept) {
    const uint32_t k = *k_ptr;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 g = (u64)blockIdx.x * blockDim.x + threadIdx.x; g < k; g += stride) {
        const uint32_t s = uniq_session[g];
        const bool known = s < n_sessions;
        u64 lat = known ? latest[s] : 0, m = known ? mask[s] : 0;
        bool st = known ? started[s] != 0 : false;
        for (u64 j = csr[g]; j < csr[g + 1]; j++) {
            const uint32_t i = perm[j];
            const u64 q = seq48[i];
            bool ok = known && q <= kSeq48Max;
            if (ok && st && q <= lat) {
                if (lat >= 64 + q) ok = false;                          // older than the window
                else if ((m >> (lat - q)) & 1) ok = false;              // already seen
            }
            if (ok && (!authentic || authentic[i])) {
                if (!st) { st = true; lat = q; m = 1; }
                else if (q > lat) { const u64 sh = q - lat; m = sh >= 64 ? 0 : m << sh; lat = q; m |= 1; }
                else m |= 1ull << (lat - q);
                accept[i] = 1;
            } else accept[i] = 0;
        }
        if (known) { latest[s] = lat; mask[s] = m; started[s] = st ? 1 : 0; }
    }
}

// ---- sequence numbers of outgoing records ---------------------------------------------------------------
// The records were grouped by session with the stable demux (arrival order kept inside a session): the record at
// sorted position j of group g is the (j - csr[g])-th of its session in this batch.  One thread per position; its group
// by binary search over csr (k + 1 entries, L2-resident).
__global__ void __launch_bounds__(256) seq_assign_kernel(u64 n, const uint32_t* __restrict__ k_ptr, const uint32_t* __restrict__ uniq_session,
                                                        const u64* __restrict__ csr, const uint32_t* __restrict__ perm,
                                                        uint32_t n_sessions, const u64* __restrict__ next_seq,
                                                        u64* __restrict__ seq48, uint8_t* __restrict__ ok) {
    const uint32_t k = *k_ptr;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
        uint32_t lo = 0, hi = k;                                         // csr[lo] <= j < csr[hi]
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (csr[mid] <= j) lo = mid; else hi = mid;
        }
        const uint32_t s = uniq_session[lo], i = perm[j];
        u64 q = 0;
        bool good = s < n_sessions;
        if (good) {
            const u64 base = next_seq[s], r = j - csr[lo];
            good = base <= kSeq48Max && r <= kSeq48Max - base;           // ErrSequenceNumberOverflow beyond 2^48 - 1
            if (good) q = base + r;
        }
        seq48[i] = q;
        ok[i] = good ? 1 : 0;
    }
}

// after the numbers are out: every session's counter advances by its records, saturating one past the last number
__global__ void __launch_bounds__(256) seq_advance_kernel(const uint32_t* __restrict__ k_ptr, const uint32_t* __restrict__ uniq_session,
                                                         const u64* __restrict__ csr, uint32_t n_sessions, u64* __restrict__ next_seq) {
    const uint32_t k = *k_ptr;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 g = (u64)blockIdx.x * blockDim.x + threadIdx.x; g < k; g += stride) {
        const uint32_t s = uniq_session[g];
        if (s >= n_sessions) continue;
        const u64 base = next_seq[s], cnt = csr[g + 1] - csr[g];
        next_seq[s] = (base > kSeq48Max || cnt > kSeq48Max + 1 - base) ? kSeq48Max + 1 : base + cnt;
    }
}

}  // namespace p2pt
