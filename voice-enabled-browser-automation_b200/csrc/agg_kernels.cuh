// agg_kernels.cuh -- device-side capture aggregator (App.tsx:148-157, :273-288), sm_100a.
//
// The reference appends 128-sample quanta to a per-session buffer and, when a quantum arrives with
// now - lastSend >= desiredMs, converts the whole buffer and resets it.  Round 1 ran that per-session
// state machine on host threads (5.2 of 5.7 ms per 65 536-session push).  Here the whole push is
// planned on the GPU; the host only enqueues a fixed sequence of kernels and reads three numbers back.
//
// One push of n quanta (quantum i belongs to sessions[i], arrived at arrival[i]):
//   1 k_agg_count    count[s]++ per quantum; the first toucher of a session appends it to `touched`
//   2 k_agg_reserve  where each touched session's index list starts (one atomic per block: the lists may sit in any order)
//                    (steps 2 and 3 return at once when no session has two quanta in the push: AggCtl::any_multi)
//   3 k_agg_scatter  sorted[cursor[s]++] = i   (atomic, so the order inside a session is arbitrary ...)
//   4 k_agg_walk     ... one thread per touched session sorts its list by i (arrival order) and runs the
//                    reference's rule over it: fill++, flush when arrival - lastSend >= desiredMs.  It only
//                    writes scratch (the session's new state, per-quantum flush marks and staging slots).
//   5 scan           over quanta: frame index and output offset of every flush, in flush order
//                    (= ascending i); its apply pass also emits the caller's frame arrays.
//   6 k_agg_convert_one / _pipe / _warp   every flushed frame: the quanta carried over from earlier pushes were
//                    picked and encoded when they arrived (step 8) and are a ready int16 prefix in
//                    staged[session][] -- copied; this push's quanta are brought into shared memory with
//                    async copies straight from the push, picked, encoded, stored; with the meter on,
//                    each of those quanta's rms is computed from the shared-memory copy.
//   7 k_agg_levels   level = max(rms, level * 0.85) along each session's flushed quanta, in order
//   8 k_agg_append   quanta that did not flush are converted NOW: the outputs whose pick falls inside the
//                    quantum go to staged[session][out_lo[slot] .. out_lo[slot + 1]) (and its rms is stored)
//   9 (commit, inside the append kernel)   fill / lastSend take their new values; per-push counters are cleared
// Steps 6-9 look at the push's status word first: if any check failed (unknown session, a buffer
// longer than max_frame_quanta, too many frames, output too small) they change nothing, which IS the
// rollback -- the plan so far lives only in scratch.
#pragma once
#include "pcm_kernels.cuh"

namespace pcmb200 {

enum : uint32_t {
    AGG_ST_BAD_SESSION = 1u,     // a session id >= max_sessions
    AGG_ST_FRAME_TOO_LONG = 2u,  // a session's buffer would exceed max_frame_quanta
    AGG_ST_TOO_MANY_FRAMES = 4u, // more flushed frames than max_frames
    AGG_ST_OUT_TOO_SMALL = 8u    // flushed frames exceed the output capacity
};

struct AggCtl {                  // lives in device memory; copied to the host after every push
    uint32_t status;
    uint32_t n_touched;
    uint32_t n_frames;
    uint32_t any_multi;          // some session has more than one quantum in this push (set by k_agg_count, one atomic per block).
                                 // Clear in the streaming regime: every index list is its session's single quantum, and the
                                 // reserve and scatter steps have nothing to do (30 of a 2^20-session push's 290 us)
    unsigned long long out_total;   // output samples incl. the padding of every frame to 8 samples
    uint32_t n_listed;           // running total of the per-session index lists (k_agg_reserve)
    uint32_t pad_;
};

struct AggFrameDesc {
    uint32_t session;
    uint32_t nq;                 // quanta in the frame
    uint32_t ring_count;         // of which the first ring_count were waiting in staged[session] (the name predates the int16 staging)
    uint32_t pos;                // the rest: sorted[pos], sorted[pos + 1], ...
};

struct AggGeom {
    uint32_t max_sessions, quantum, half_quanta;
    uint32_t stride;             // integer ratio R, or 0 when picks come from the table
    double desired_ms;
};

// Programmatic dependent launch (opt-out: tune agg_pdl = 0).  A push is a chain of eight or nine short kernels, each
// needing everything its predecessor wrote; launched the plain way, every link pays a full launch latency after
// the previous grid has drained.  With the programmatic-serialisation launch attribute a grid may be scheduled
// as soon as every block of its predecessor has passed pdl_enter()'s trigger (or exited), and its blocks then
// sit in `griddepcontrol.wait` until the predecessor has completed and its writes are visible.  EVERY block of
// EVERY kernel in the chain calls pdl_enter() first thing, unconditionally: a grid therefore cannot complete
// before its predecessor has, which is what makes results of kernels further up the chain safe to read too.
// Launched without the attribute both instructions do nothing.
// With the grids of a push overlapping like that, an SM's L1 is no longer wiped between two kernels of the chain: a
// line that a block of kernel K brought in (say the control block, read at K's start) can still sit there when a
// block of kernel K+1 reads it on the same SM, after K's last block has changed it in L2.  Seen on the B200: a push
// refused for lack of output space was nevertheless committed by some blocks, which read a stale status word.  So
// everything that another kernel of the push (or the push before) has written is loaded with ld.global.cg -- L2 is
// the point of coherence -- and only true inputs (the caller's arrays, the pick table) go through L1.
template <typename T> __device__ __forceinline__ T ld_l2(const T *p) { return __ldcg(p); }
__device__ __forceinline__ AggFrameDesc ld_l2(const AggFrameDesc *p)
{
    const uint4 r = __ldcg(reinterpret_cast<const uint4 *>(p));
    return AggFrameDesc{r.x, r.y, r.z, r.w};
}

// The head of the push's control block, fetched ONCE per block (thread 0, through L2) and handed to every thread:
// a million threads asking L2 for the same word cost several microseconds per kernel.  Every thread of the block
// must call it (it holds a barrier), right after pdl_enter().
struct CtlHead { uint32_t status, n_touched, n_frames, any_multi; };
__device__ __forceinline__ CtlHead ctl_head(const AggCtl *ctl)
{
    __shared__ uint4 head;
    if (threadIdx.x == 0) head = __ldcg(reinterpret_cast<const uint4 *>(ctl));   // {status, n_touched, n_frames, any_multi}
    __syncthreads();
    return CtlHead{head.x, head.y, head.z, head.w};
}

__device__ __forceinline__ void pdl_enter()
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// Do two neighbouring lanes of this warp hold the same (valid) session?  All 32 lanes must call.
__device__ __forceinline__ bool warp_has_neighbours(uint32_t s, uint32_t lane)
{
    const uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, s, 1);
    return __any_sync(0xFFFFFFFFu, lane != 0u && prev == s && s != 0xFFFFFFFFu) != 0;
}

// ---- 1: count ------------------------------------------------------------------------------------------
// Exclusive scan of one value per thread over a 256-thread block (all threads must call); `total` = the block's sum.
__device__ __forceinline__ uint32_t block_exscan256(uint32_t v, uint32_t *warp_sums, uint32_t &total)
{
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= (uint32_t)d) inc += o;
    }
    if (lane == 31u) warp_sums[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const uint32_t ws = warp_sums[w];
        if ((uint32_t)w < warp) base += ws;
        tot += ws;
    }
    __syncthreads();
    total = tot;
    return base + inc - v;
}

// Space for `v` items per thread out of one running total: ONE atomic per block (a million threads -- or even their
// 32 768 warps -- adding to the same address serialise in L2 at about a nanosecond each: 33 us per kernel, measured).
__device__ __forceinline__ uint32_t block_reserve256(uint32_t v, uint32_t *counter, uint32_t *smem9)
{
    uint32_t total;
    const uint32_t off = block_exscan256(v, smem9, total);
    if (threadIdx.x == 0) smem9[8] = total ? atomicAdd(counter, total) : 0u;
    __syncthreads();
    return smem9[8] + off;
}

__global__ void __launch_bounds__(256)
k_agg_count(const uint32_t *__restrict__ sessions, uint32_t n, uint32_t max_sessions, uint32_t *__restrict__ count,
            uint32_t *__restrict__ touched, uint32_t *__restrict__ touched_idx, AggCtl *__restrict__ ctl)
{
    pdl_enter();
    __shared__ uint32_t sm[9];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t s = 0xFFFFFFFFu;
    bool first = false, again = false;
    if (i < n) {
        s = __ldg(sessions + i);
        if (s >= max_sessions) { atomicOr(&ctl->status, AGG_ST_BAD_SESSION); s = 0xFFFFFFFFu; }
    }
    {
        // Lanes of a warp that hold quanta of the SAME session (a push grouped by session: 23 neighbours) add to its
        // counter once, through one of them: 32 atomics on one address from one warp serialise (count 25 us instead of
        // 11 per 1.5 M quanta, measured).  Finding the groups (match.any) costs 9 us per 1.5 M quanta when every lane
        // holds a different session, so it is done only in warps where two neighbouring lanes agree.
        const uint32_t lane = threadIdx.x & 31u;
        if (warp_has_neighbours(s, lane)) {
            const unsigned peers = __match_any_sync(0xFFFFFFFFu, s);
            const uint32_t leader = (uint32_t)__ffs((int)peers) - 1u;
            uint32_t old = 0u;
            if (lane == leader && s != 0xFFFFFFFFu) old = atomicAdd(count + s, (uint32_t)__popc(peers));
            old = __shfl_sync(0xFFFFFFFFu, old, (int)leader);
            if (s != 0xFFFFFFFFu) {
                first = old == 0u && lane == leader;                   // the first toucher lists the session ...
                again = !first;
            }
        } else if (s != 0xFFFFFFFFu) {
            first = atomicAdd(count + s, 1u) == 0u;
            again = !first;
        }
    }
    const uint32_t at = block_reserve256(first ? 1u : 0u, &ctl->n_touched, sm);
    if (first) { touched[at] = s; touched_idx[at] = i; }               // ... and its own index: the whole list if nobody follows
    if (__syncthreads_or(again) && threadIdx.x == 0) atomicOr(&ctl->any_multi, 1u);
}

// ---- 2: reserve --------------------------------------------------------------------------------------------
// Where each touched session's index list starts in `sorted`.  The lists may sit in any order, so no grid-wide
// scan is needed: every block takes the slots of its sessions from a running total.
__global__ void __launch_bounds__(256)
k_agg_reserve(const uint32_t *__restrict__ touched, const uint32_t *__restrict__ count, uint32_t *__restrict__ cursor,
              AggCtl *__restrict__ ctl)
{
    pdl_enter();
    __shared__ uint32_t sm[9];
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    const CtlHead head = ctl_head(ctl);
    if (!head.any_multi) return;                                       // single-quantum lists: the walk kernel needs no cursor
    const bool live = k < head.n_touched;
    const uint32_t s = live ? ld_l2(touched + k) : 0u, c = live ? ld_l2(count + s) : 0u;
    const uint32_t at = block_reserve256(c, &ctl->n_listed, sm);       // one atomic per block
    if (live) cursor[s] = at;
}

// ---- 5: scan over the push's quanta (two channels, 2048 items per block) ---------------------------------------
// Frame number and output offset of every flush, in flush order (= ascending quantum index): a two-pass
// scan -- per-block totals, then per-block rescan + emit (each block adds up the totals before it).  Channel a counts the
// flushes, channel b adds up the frames' output lengths rounded up to 8 samples.
constexpr int kScanThreads = 256, kScanItems = 8, kScanBlock = kScanThreads * kScanItems;

struct ScanArgs {
    uint32_t n;                      // item i < n: a = fl_nq[i] != 0, b = padded output length of that frame
    const uint32_t *fl_nq, *nout_of;
    const AggCtl *ctl;
};

// The kScanItems consecutive items of one thread: va[k] = 1 if quantum idx0 + k flushes a frame, vb[k] = that
// frame's output length rounded up to 8 samples.  The flags are read as two 128-bit loads (the per-item form was a
// chain of three dependent loads per item and made these three passes latency-bound).
__device__ __forceinline__ void scan_load8(const ScanArgs &a, uint32_t status, uint32_t idx0, uint32_t len, uint32_t (&va)[kScanItems], uint32_t (&vb)[kScanItems])
{
    static_assert(kScanItems == 8, "two uint4 loads per thread");
    uint32_t nq[kScanItems];
#pragma unroll
    for (int k = 0; k < kScanItems; k++) nq[k] = 0u;
    // a rejected push leaves fl_nq partly unwritten (the walk stops at the first problem): read nothing
    if (status == 0u && idx0 < len) {
        if (idx0 + kScanItems <= len) {
            const uint4 lo = __ldcg(reinterpret_cast<const uint4 *>(a.fl_nq + idx0)), hi = __ldcg(reinterpret_cast<const uint4 *>(a.fl_nq + idx0 + 4));
            nq[0] = lo.x; nq[1] = lo.y; nq[2] = lo.z; nq[3] = lo.w; nq[4] = hi.x; nq[5] = hi.y; nq[6] = hi.z; nq[7] = hi.w;
        } else {
#pragma unroll
            for (int k = 0; k < kScanItems; k++) if (idx0 + k < len) nq[k] = ld_l2(a.fl_nq + idx0 + k);
        }
    }
#pragma unroll
    for (int k = 0; k < kScanItems; k++) {
        va[k] = nq[k] ? 1u : 0u;
        vb[k] = nq[k] ? (__ldg(a.nout_of + nq[k]) + 7u) & ~7u : 0u;
    }
}

// exclusive block scan of one value per thread (256 threads); returns the block total in `total`
__device__ __forceinline__ uint32_t block_exscan(uint32_t v, uint32_t *warp_sums, uint32_t &total)
{
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= (uint32_t)d) inc += o;
    }
    if (lane == 31u) warp_sums[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; w++) {
        const uint32_t ws = warp_sums[w];
        if ((uint32_t)w < warp) base += ws;
        tot += ws;
    }
    __syncthreads();
    total = tot;
    return base + inc - v;
}

__global__ void __launch_bounds__(kScanThreads)
k_scan_reduce(ScanArgs a, uint32_t *__restrict__ sums_a, uint32_t *__restrict__ sums_b)
{
    pdl_enter();
    __shared__ uint32_t ws[kScanThreads / 32];
    const uint32_t status = ctl_head(a.ctl).status;
    const uint32_t len = a.n;
    const uint32_t base = blockIdx.x * kScanBlock;
    if (base >= len) return;
    uint32_t va[kScanItems], vb[kScanItems], sa = 0, sb = 0;
    scan_load8(a, status, base + threadIdx.x * kScanItems, len, va, vb);
#pragma unroll
    for (int k = 0; k < kScanItems; k++) { sa += va[k]; sb += vb[k]; }
    uint32_t ta, tb;
    block_exscan(sa, ws, ta);
    block_exscan(sb, ws, tb);
    if (threadIdx.x == 0) { sums_a[blockIdx.x] = ta; sums_b[blockIdx.x] = tb; }
}

struct ScanApplyB {
    const uint32_t *sessions, *fl_ring, *fl_pos;
    uint32_t *frame_session, *frame_len;
    unsigned long long *frame_offset;
    AggFrameDesc *fdesc;
    unsigned long long *fdesc_off;
    uint32_t max_frames;
};

// Second pass: every block adds up the totals of the blocks before it (a few hundred numbers: cheaper than a third
// kernel whose only job was to scan them), rescans its own items and emits the frames.  The last block knows the
// grand totals: it publishes n_frames / out_total and checks the two capacities.  Frames emitted by earlier blocks
// of a push that then turns out too large are harmless: nothing is written past max_frames, and nothing
// downstream runs once the status word is set.
__global__ void __launch_bounds__(kScanThreads, 5)         // (a push of 1.5 M quanta is 736 blocks: one wave at five per SM)
k_scan_apply(ScanArgs a, const uint32_t *__restrict__ sums_a, const uint32_t *__restrict__ sums_b, ScanApplyB ap,
             AggCtl *__restrict__ ctl, unsigned long long out_capacity)
{
    pdl_enter();
    __shared__ uint32_t ws[kScanThreads / 32];
    __shared__ unsigned long long red_a[kScanThreads / 32], red_b[kScanThreads / 32];
    const uint32_t status = ctl_head(a.ctl).status;
    const uint32_t len = a.n;
    const uint32_t base = blockIdx.x * kScanBlock;
    if (base >= len) return;
    if (status != 0u) return;                          // rejected push: nothing to emit
    unsigned long long pa = 0, pb = 0;                 // totals of the blocks before this one
    for (uint32_t b = threadIdx.x; b < blockIdx.x; b += kScanThreads) { pa += ld_l2(sums_a + b); pb += ld_l2(sums_b + b); }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { pa += __shfl_xor_sync(0xFFFFFFFFu, pa, d); pb += __shfl_xor_sync(0xFFFFFFFFu, pb, d); }
    if ((threadIdx.x & 31u) == 0u) { red_a[threadIdx.x >> 5] = pa; red_b[threadIdx.x >> 5] = pb; }
    uint32_t va[kScanItems], vb[kScanItems], sa = 0, sb = 0;
    scan_load8(a, status, base + threadIdx.x * kScanItems, len, va, vb);
#pragma unroll
    for (int k = 0; k < kScanItems; k++) { sa += va[k]; sb += vb[k]; }
    uint32_t ta, tb;
    const uint32_t xa = block_exscan(sa, ws, ta);      // (its barriers also publish red_a / red_b)
    const uint32_t xb = block_exscan(sb, ws, tb);
    pa = 0; pb = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; w++) { pa += red_a[w]; pb += red_b[w]; }
    // The block's frames are first listed in shared memory (quantum index and output offset within the block, in frame
    // order), then emitted with thread <-> frame: the stores to the five per-frame arrays are then contiguous across a
    // warp.  Emitted straight from the scan, a thread wrote its (up to) eight frames one after the other, every store 32
    // sectors wide and every frame's loads queued behind the previous frame's stores: 54 us for 2^20 frames.
    __shared__ uint32_t st_idx[kScanBlock], st_off[kScanBlock];
    {
        uint32_t la = xa, lb = xb;
#pragma unroll
        for (int k = 0; k < kScanItems; k++) {
            if (va[k]) { st_idx[la] = base + threadIdx.x * kScanItems + k; st_off[la] = lb; }
            la += va[k]; lb += vb[k];
        }
    }
    __syncthreads();
    // four frames per trip, everything they need asked for before anything is stored (the per-frame arrays may alias
    // as far as the compiler knows, so a frame's loads would otherwise wait behind the previous frame's stores)
    for (uint32_t r0 = threadIdx.x; r0 < ta; r0 += 4u * kScanThreads) {
        uint32_t e_nq[4], e_s[4], e_rc[4], e_pos[4], e_len[4];
#pragma unroll
        for (uint32_t u = 0; u < 4u; u++) {
            const uint32_t r = r0 + u * kScanThreads;
            e_nq[u] = e_s[u] = e_rc[u] = e_pos[u] = 0u;
            if (r < ta) {
                const uint32_t idx = st_idx[r];
                e_nq[u] = ld_l2(a.fl_nq + idx); e_s[u] = ap.sessions[idx];
                e_rc[u] = ld_l2(ap.fl_ring + idx); e_pos[u] = ld_l2(ap.fl_pos + idx);
            }
        }
#pragma unroll
        for (uint32_t u = 0; u < 4u; u++) e_len[u] = __ldg(a.nout_of + e_nq[u]);
#pragma unroll
        for (uint32_t u = 0; u < 4u; u++) {
            const uint32_t r = r0 + u * kScanThreads, f = (uint32_t)pa + r;  // quantum st_idx[r] flushes frame number pa + r
            if (r < ta && f < ap.max_frames) {
                const unsigned long long off = pb + st_off[r];
                if (ap.frame_session) ap.frame_session[f] = e_s[u];
                if (ap.frame_len) ap.frame_len[f] = e_len[u];
                if (ap.frame_offset) ap.frame_offset[f] = off;
                ap.fdesc[f] = AggFrameDesc{e_s[u], e_nq[u], e_rc[u], e_pos[u]};
                ap.fdesc_off[f] = off;
            }
        }
    }
    if (blockIdx.x == gridDim.x - 1u && threadIdx.x == 0u) {
        const unsigned long long total_a = pa + ta, total_b = pb + tb;
        ctl->n_frames = (uint32_t)total_a;
        ctl->out_total = total_b;
        uint32_t st = 0;
        if (total_a > ap.max_frames) st |= AGG_ST_TOO_MANY_FRAMES;
        if (total_b > out_capacity) st |= AGG_ST_OUT_TOO_SMALL;
        if (st) atomicOr(&ctl->status, st);
    }
}

// ---- 3: scatter --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_agg_scatter(const uint32_t *__restrict__ sessions, uint32_t n, uint32_t *__restrict__ cursor, uint32_t *__restrict__ sorted,
              const AggCtl *__restrict__ ctl)
{
    pdl_enter();
    const CtlHead head = ctl_head(ctl);
    if (head.status != 0u || !head.any_multi) return;                  // (single-quantum lists are written by the walk kernel)
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31u;
    const uint32_t s = i < n ? __ldg(sessions + i) : 0xFFFFFFFFu;
    // one atomic per session per warp where neighbouring lanes share a session, as in the count step
    if (warp_has_neighbours(s, lane)) {
        const unsigned peers = __match_any_sync(0xFFFFFFFFu, s);
        const uint32_t leader = (uint32_t)__ffs((int)peers) - 1u;
        uint32_t at = 0u;
        if (lane == leader && i < n) at = atomicAdd(cursor + s, (uint32_t)__popc(peers));
        at = __shfl_sync(0xFFFFFFFFu, at, (int)leader);
        if (i < n) sorted[at + (uint32_t)__popc(peers & ((1u << lane) - 1u))] = i;
    } else if (i < n) {
        sorted[atomicAdd(cursor + s, 1u)] = i;
    }
}

// ---- 4: walk -------------------------------------------------------------------------------------------------
// In-place heap sort for the rare session with a long list in one push (short lists use insertion sort).
// (the list was written by the scatter kernel: its elements are read through L2, see ld_l2; the thread's own stores
//  are seen by its own later loads in program order)
__device__ inline void sift_down(uint32_t *a, uint32_t start, uint32_t end)
{
    uint32_t root = start;
    while (2u * root + 1u < end) {
        uint32_t child = 2u * root + 1u;
        if (child + 1u < end && ld_l2(a + child) < ld_l2(a + child + 1u)) child++;
        const uint32_t r = ld_l2(a + root), c = ld_l2(a + child);
        if (r >= c) return;
        a[root] = c; a[child] = r;
        root = child;
    }
}
__device__ inline void sort_indices(uint32_t *a, uint32_t m)
{
    if (m <= 48u) {
        for (uint32_t i = 1; i < m; i++) {
            const uint32_t v = ld_l2(a + i);
            uint32_t j = i;
            while (j > 0) {
                const uint32_t p = ld_l2(a + j - 1);
                if (p <= v) break;
                a[j] = p; j--;
            }
            if (j != i) a[j] = v;
        }
        return;
    }
    for (uint32_t s = m / 2; s-- > 0;) sift_down(a, s, m);
    for (uint32_t e = m; e-- > 1;) {
        const uint32_t t = ld_l2(a), u = ld_l2(a + e);
        a[0] = u; a[e] = t;
        sift_down(a, 0, e);
    }
}

struct AggWalkOut {
    uint32_t *fl_nq, *fl_ring, *fl_pos;   // [n] per quantum: frame length in quanta if it flushes (else 0), ...
    uint32_t *dst_slot;                   // [n] staging slot (session * half_quanta*2 + position) of a quantum that does not flush, else 0xFFFFFFFF
    uint32_t *new_fill, *flushed;         // [touched] new fill; push quanta of the session that flushed
    double *new_last;                     // [touched] new lastSend
};

// MREG: the longest index list kept in registers (longer ones are sorted in global memory).  32 covers a 60 ms frame
// arriving in one push, at 96 registers per thread -- 640 threads per SM, which is what a push of 2^20 sessions with
// ONE quantum each then pays for: 51 us, against 20 for the MREG = 4 build of the same code.  The host picks the
// build from the shape of the previous push (quanta per touched session); either is correct for any push.
template <uint32_t MREG>
__global__ void __launch_bounds__(128)
k_agg_walk(AggGeom g, const double *__restrict__ arrival, const uint32_t *__restrict__ touched, const uint32_t *__restrict__ touched_idx,
           const uint32_t *__restrict__ count, uint32_t *__restrict__ cursor, uint32_t *__restrict__ sorted, const uint32_t *__restrict__ fill,
           const double *__restrict__ last_send, AggWalkOut o, AggCtl *__restrict__ ctl)
{
    pdl_enter();
    const CtlHead head = ctl_head(ctl);
    if (head.status != 0u) return;
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= head.n_touched) return;
    // No session with two quanta in this push (the streaming regime): the list of touched session k is the index its
    // one quantum left in touched_idx, it goes to sorted[k], and cursor gets the value the scatter step would have left
    // (one past the list) for the level kernel.
    const bool single = head.any_multi == 0u;
    const uint32_t s = ld_l2(touched + k);
    uint32_t m = 1u, base = k, only = 0u;
    if (single) {
        only = ld_l2(touched_idx + k);
        sorted[k] = only;
        cursor[s] = k + 1u;
    } else {
        m = ld_l2(count + s);
        base = ld_l2(cursor + s) - m;
    }
    uint32_t *list = sorted + base;
    const uint32_t c0 = ld_l2(fill + s);
    uint32_t f = c0, ring_cnt = c0, frame_start = 0;
    double last = ld_l2(last_send + s);
    bool too_long = false;
    // the reference's rule for the j-th quantum of the session in this push (push index i, arrival time `now`)
    auto rule = [&](uint32_t i, double now, uint32_t j) {
        if (f >= g.half_quanta) { too_long = true; return; }
        f++;
        // App.tsx:279-280: elapsed = performance.now() - lastSend; if (elapsed >= desiredMs) flush
        if (now - last >= g.desired_ms) {
            o.fl_nq[i] = f; o.fl_ring[i] = ring_cnt; o.fl_pos[i] = base + frame_start;
            f = 0; ring_cnt = 0; frame_start = j + 1u;
            last = now;                                                        // App.tsx:287
        } else {
            o.fl_nq[i] = 0u;
        }
    };
    // The rule is sequential in `last`, but the addresses are not: arrival times are fetched in batches of 8 (one
    // memory round trip per batch instead of one per quantum -- the kernel was a chain of dependent misses).
    if (single) {                                                              // one quantum, one rule, done
        rule(only, arrival[only], 0u);
        if (too_long) { atomicOr(&ctl->status, AGG_ST_FRAME_TOO_LONG); return; }
        o.dst_slot[only] = frame_start == 0u ? s * g.half_quanta + c0 : 0xFFFFFFFFu;
        o.new_fill[k] = f;
        o.new_last[k] = last;
        o.flushed[k] = frame_start;
        return;
    }
    constexpr uint32_t B = MREG < 8u ? MREG : 8u;
    if (m <= MREG) {
        // The usual case (~23 quanta per 60 ms, 1 per push when streaming): the index list lives in registers.
        // Arrival order by a bitonic network -- 240 compare-exchanges, no memory traffic, no divergence; the
        // in-place insertion sort this replaces stored to the list on every step and then waited for its own
        // stores to come back from L2 (a third of the kernel, ncu).
        uint32_t v[MREG];
#pragma unroll
        for (int u = 0; u < (int)MREG; u++) v[u] = (uint32_t)u < m ? ld_l2(list + u) : 0xFFFFFFFFu;
        if (m > 1u) {
#pragma unroll
            for (int kk = 2; kk <= (int)MREG; kk <<= 1) {
#pragma unroll
                for (int jj = kk >> 1; jj > 0; jj >>= 1) {
#pragma unroll
                    for (int i = 0; i < (int)MREG; i++) {
                        const int l = i ^ jj;
                        if (l > i) {
                            const uint32_t lo = min(v[i], v[l]), hi = max(v[i], v[l]);
                            const bool up = (i & kk) == 0;
                            v[i] = up ? lo : hi;
                            v[l] = up ? hi : lo;
                        }
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < (int)MREG; u++) if ((uint32_t)u < m) list[u] = v[u];  // the convert and level kernels read it
        }
#pragma unroll
        for (int b0 = 0; b0 < (int)MREG; b0 += (int)B) {
            if ((uint32_t)b0 < m) {
                double at[B];
#pragma unroll
                for (int u = 0; u < (int)B; u++) at[u] = (uint32_t)(b0 + u) < m ? arrival[v[b0 + u]] : 0.0;
#pragma unroll
                for (int u = 0; u < (int)B; u++) if ((uint32_t)(b0 + u) < m) rule(v[b0 + u], at[u], (uint32_t)(b0 + u));
            }
        }
        if (too_long) { atomicOr(&ctl->status, AGG_ST_FRAME_TOO_LONG); return; }
        // quanta after the last flush stay in the session's buffer: behind what was there if nothing flushed
        const uint32_t first_slot = s * g.half_quanta + (frame_start == 0u ? c0 : 0u);
#pragma unroll
        for (int u = 0; u < (int)MREG; u++)
            if ((uint32_t)u < m) o.dst_slot[v[u]] = (uint32_t)u >= frame_start ? first_slot + ((uint32_t)u - frame_start) : 0xFFFFFFFFu;
    } else {
        sort_indices(list, m);
        for (uint32_t j0 = 0; j0 < m; j0 += B) {
            uint32_t idx[B];
            double at[B];
#pragma unroll
            for (uint32_t u = 0; u < B; u++) idx[u] = j0 + u < m ? ld_l2(list + j0 + u) : 0u;
#pragma unroll
            for (uint32_t u = 0; u < B; u++) at[u] = j0 + u < m ? arrival[idx[u]] : 0.0;
#pragma unroll
            for (uint32_t u = 0; u < B; u++) if (j0 + u < m) rule(idx[u], at[u], j0 + u);
        }
        if (too_long) { atomicOr(&ctl->status, AGG_ST_FRAME_TOO_LONG); return; }
        const uint32_t first_slot = s * g.half_quanta + (frame_start == 0u ? c0 : 0u);
        for (uint32_t j = 0; j < m; j++) o.dst_slot[ld_l2(list + j)] = j >= frame_start ? first_slot + (j - frame_start) : 0xFFFFFFFFu;
    }
    o.new_fill[k] = f;
    o.new_last[k] = last;
    o.flushed[k] = frame_start;
}

// 16-byte asynchronous global -> shared copy (LDGSTS, L2 only).  The aggregator stages single 512-byte
// quanta that are scattered through the push; one bulk copy per quantum turned out to be bound by the
// copy engine's per-request cost (~80 clocks per request, measured), so the quanta are moved as plain
// 16-byte async copies issued by all threads: a warp-wide instruction is one contiguous 512-byte quantum.
__device__ __forceinline__ void cp_async_16(void *dst_smem, const void *src_gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src_gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.commit_group;\n cp.async.wait_group 0;" ::: "memory");
}

// ---- 6: convert -------------------------------------------------------------------------------------------------
// A CTA takes G consecutive frames per step.  Their quanta are packed into `slots` of shared memory (slot pitch =
// quantum + 4 floats) with 16-byte async copies, all in flight together: a warp-wide copy instruction moves one
// 512-byte quantum.  Then all threads pick + encode + store, one PAIR of adjacent outputs per lane (lane stride
// 2 * ratio words in shared memory: at most 2-way bank conflicts; lane <-> 8 outputs was 8-way for ratio 3), one
// 32-bit streaming store per lane, 128 contiguous bytes per warp.  With METER the first T lanes each accumulate
// their quantum's sum of squares in sample order (same arithmetic as kernel 2c).
// Two forms: k_agg_convert_one (G = 1, at most 64 quanta per frame, at most 128 threads: the usual ~60 ms frame)
// and k_agg_convert_pipe (everything else).  How they got here, by ncu (65 536 frames of 23 quanta, 0.9 GB):
//   step by step, one barrier per dependent round trip, lane <-> 8 outputs           200 us, issue slots 78 % busy
//   + next step's descriptors fetched while this step's copies fly (pipe)            200 us: it was issue-bound --
//       170 M warp-instructions, half of them a generic copy loop that spent ~70 instructions per quantum
//   + copy loop specialised for the 128-sample quantum, lean pair loop               160 us, 103 M instructions
//   + descriptors in registers, prefetched two frames ahead, persistent grid (one)   ~150 us, 0.92 of the HBM peak
struct AggConvertArgs {
    const float *samples;            // the push's quanta
    const int16_t *staged;           // [max_sessions][nout_pad] outputs of the quanta that were waiting (see AggAppendArgs)
    const uint32_t *out_lo;          // [half_quanta + 1] first output of quantum j
    uint32_t nout_pad;
    const uint32_t *sorted;
    const AggFrameDesc *fdesc;
    const unsigned long long *fdesc_off;
    const uint32_t *nout_of;         // [half_quanta + 1] output length of a frame of nq quanta
    const uint32_t *lut;             // pick table shared by every frame length (prefix property), or null
    int16_t *out;
    double *rms_at;                  // [n] rms of push quantum sorted[p] at rms_at[p] (METER)
    const AggCtl *ctl;
    uint32_t G;                      // frames per CTA step
    uint32_t magic_q;                // ceil(2^32 / quantum); never 0 here (quantum >= 4)
    int out_aligned16;               // `out` is 16-byte aligned: 32-bit stores (else 16-bit ones)
    int hint;                        // unused by the convert kernels (their stores are always streaming)
};

// ---- 6 (general form): convert, software-pipelined ----------------------------------------------------------------
// What ncu said about the first, step-by-step version shaped this one.  There a CTA step was a chain of
// dependent round trips -- frame descriptor (L2) -> index list entry (L2) -> the quanta themselves (HBM) -> picks ->
// stores -- with a block barrier between each, and the arena sat empty during the first two: the bytes a CTA keeps
// in flight are what its shared memory buys, and for half of every step it bought nothing.  Here
//   * the descriptors and source addresses of step i+1 are fetched while the async copies of step i are in flight
//     (two meta buffers; warp 0 turns the step's descriptors into slot / pair offsets with one warp scan), so the
//     arena goes from "consumed" straight back to "filling";
//   * a step has 3 block barriers instead of 5;
//   * the picks are lane <-> PAIR of adjacent outputs (lane stride 2*ratio words in shared memory: at most 2-way bank
//     conflicts; lane <-> 8 outputs was 8-way for ratio 3), written as one 32-bit streaming store per lane, 128
//     contiguous bytes per warp.
struct AggMetaLayout { uint32_t slot_src, foff, fd, qstart, pstart, nout, bytes; };
__host__ __device__ inline AggMetaLayout agg_meta_layout(uint32_t G, uint32_t max_slots)
{
    AggMetaLayout L;
    uint32_t o = 0;
    L.slot_src = o; o += max_slots * 8u;       // const float * per slot
    L.foff = o; o += G * 8u;                   // output offset of each frame
    L.fd = o; o += G * 16u;                    // AggFrameDesc
    L.qstart = o; o += (G + 1u) * 4u;          // first slot of each frame
    L.pstart = o; o += (G + 1u) * 4u;          // first output pair of each frame
    L.nout = o; o += G * 4u;                   // output length of each frame
    L.bytes = (o + 15u) & ~15u;
    return L;
}
__host__ __device__ inline size_t agg_arena_bytes(uint32_t G, uint32_t half_quanta, uint32_t quantum)
{
    return (((size_t)G * half_quanta * (quantum + 4u) * 4u) + 127u) & ~(size_t)127u;
}

__device__ __forceinline__ void st_32(int16_t *p, uint32_t v, int hint)
{
    if (hint == 1) asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
    else *reinterpret_cast<uint32_t *>(p) = v;
}

struct AggMeta {
    const float **slot_src;
    unsigned long long *foff;
    AggFrameDesc *fd;
    uint32_t *qstart, *pstart, *nout;
    __device__ __forceinline__ AggMeta(unsigned char *base, const AggMetaLayout &L)
        : slot_src(reinterpret_cast<const float **>(base + L.slot_src)), foff(reinterpret_cast<unsigned long long *>(base + L.foff)),
          fd(reinterpret_cast<AggFrameDesc *>(base + L.fd)), qstart(reinterpret_cast<uint32_t *>(base + L.qstart)),
          pstart(reinterpret_cast<uint32_t *>(base + L.pstart)), nout(reinterpret_cast<uint32_t *>(base + L.nout)) {}
};

// Descriptors of frames f0 .. f0+nf-1 -> m (one block barrier inside; every thread of the CTA must call it).
__device__ __forceinline__ void agg_build_meta(const AggGeom &g, const AggConvertArgs &a, uint32_t f0, uint32_t nf, const AggMeta &m)
{
    const uint32_t lane = threadIdx.x & 31u;
    if (threadIdx.x < 32u) {                              // G <= 32: one lane per frame, offsets by a warp scan
        AggFrameDesc d{0u, 0u, 0u, 0u};
        unsigned long long off = 0;
        uint32_t no = 0;
        if (lane < nf) { d = ld_l2(a.fdesc + f0 + lane); off = ld_l2(a.fdesc_off + f0 + lane); no = __ldg(a.nout_of + d.nq); }
        uint32_t iq = d.nq, ip = (no + 1u) >> 1;
        const uint32_t vq = iq, vp = ip;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const uint32_t oq = __shfl_up_sync(0xFFFFFFFFu, iq, s), op = __shfl_up_sync(0xFFFFFFFFu, ip, s);
            if (lane >= (uint32_t)s) { iq += oq; ip += op; }
        }
        if (lane < nf) { m.fd[lane] = d; m.foff[lane] = off; m.nout[lane] = no; m.qstart[lane] = iq - vq; m.pstart[lane] = ip - vp; }
        if (lane == 31u) { m.qstart[nf] = iq; m.pstart[nf] = ip; }
    }
    __syncthreads();
    const uint32_t T = m.qstart[nf], Q = g.quantum;
    for (uint32_t sl = threadIdx.x; sl < T; sl += blockDim.x) {          // where every slot's quantum comes from
        uint32_t k = 0;
        while (k + 1u < nf && sl >= m.qstart[k + 1u]) k++;
        const uint32_t j = sl - m.qstart[k];
        const AggFrameDesc d = m.fd[k];
        // the first ring_count quanta of a frame were converted when they arrived (a.staged): nothing to fetch for them
        m.slot_src[sl] = j < d.ring_count ? nullptr : a.samples + (size_t)ld_l2(a.sorted + d.pos + (j - d.ring_count)) * Q;
    }
}

template <bool METER, bool LUT>
__global__ void __launch_bounds__(256)
k_agg_convert_pipe(AggGeom g, AggConvertArgs a)
{
    pdl_enter();
    const CtlHead head = ctl_head(a.ctl);
    if (head.status != 0u) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t Q = g.quantum, pitch = Q + 4u, gpq = Q >> 2;
    const AggMetaLayout L = agg_meta_layout(a.G, a.G * g.half_quanta);
    float *arena = reinterpret_cast<float *>(smem_raw);
    unsigned char *meta_base = smem_raw + agg_arena_bytes(a.G, g.half_quanta, Q);

    const uint32_t n_frames = head.n_frames, step = gridDim.x * a.G;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nw = blockDim.x >> 5;
    uint32_t f0 = blockIdx.x * a.G;
    if (f0 >= n_frames) return;
    agg_build_meta(g, a, f0, min(a.G, n_frames - f0), AggMeta(meta_base, L));
    __syncthreads();
    for (uint32_t it = 0;; it++) {
        const AggMeta m(meta_base + (it & 1u) * L.bytes, L);
        const uint32_t nf = min(a.G, n_frames - f0);
        const uint32_t T = m.qstart[nf];
        // all quanta of the step in flight at once: a warp-wide copy instruction moves one 512-byte quantum (the
        // slot index is warp-uniform, so no per-thread division and one broadcast read of the source address)
        if (gpq == 32u) {                                   // the AudioWorklet quantum (128 samples): the whole loop body is
            const uint32_t dst0 = smem_u32(arena) + 16u * lane;     // one LDS.64, one IMAD and the copy -- ncu showed the generic
            for (uint32_t sl = warp; sl < T; sl += nw) {    // form below at ~70 instructions per quantum, half of the kernel
                const float *src = m.slot_src[sl];          // null: that quantum was converted when it arrived
                if (src) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + sl * (pitch * 4u)), "l"(src + 4u * lane) : "memory");
            }
        } else {
            for (uint32_t sl = warp; sl < T; sl += nw) {
                const float *src = m.slot_src[sl];
                float *dst = arena + (size_t)sl * pitch;
                if (!src) continue;
#pragma unroll 1
                for (uint32_t gi = lane; gi < gpq; gi += 32u) cp_async_16(dst + 4u * gi, src + 4u * gi);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        const uint32_t f_next = f0 + step;
        const bool more = f_next < n_frames;                                   // block-uniform
        if (more) agg_build_meta(g, a, f_next, min(a.G, n_frames - f_next), AggMeta(meta_base + ((it + 1u) & 1u) * L.bytes, L));
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();

        // ---- pick + encode + store: frame by frame, work item = 2 adjacent outputs ---------------------------
        // The kernel is issue-bound (ncu: 70-80 % of the issue slots), so the inner loop carries nothing it does not
        // need: everything per frame is read once outside it, the table / stride choice is a template parameter, full
        // pairs need no tail logic, and an integer-ratio pick can never fall outside its frame (n_out = floor(n_in / R)),
        // so only table picks are range-checked.  Sample i of a frame sits at word i + 4 * (i / quantum) of its slots.
        for (uint32_t k = 0; k < nf; k++) {
            const uint32_t n_in = m.fd[k].nq * Q, n_out = m.nout[k];
            const float *fr = arena + (size_t)m.qstart[k] * pitch;
            int16_t *dstf = a.out + m.foff[k];
            const float qnan = __int_as_float(0x7FC00000);                     // outside the frame: undefined -> NaN -> 0
            // outputs below `osplit` belong to quanta that were converted when they arrived: copied from a.staged
            const uint32_t rc = m.fd[k].ring_count, osplit = rc ? __ldg(a.out_lo + rc) : 0u;
            const int16_t *st = a.staged + (size_t)m.fd[k].session * a.nout_pad;
            if (a.out_aligned16) {
                // mq = ceil(2^32 / quantum) is never 0 here (quantum >= 4), so i / quantum is one multiply-high
                const uint32_t nfull = n_out >> 1, R = g.stride, mq = a.magic_q, bd = blockDim.x;
                const uint2 *lut2 = reinterpret_cast<const uint2 *>(a.lut);
                auto pair = [&](uint32_t p) -> uint32_t {
                    if (2u * p + 1u < osplit) return ld_l2(reinterpret_cast<const uint32_t *>(st) + p);
                    if (2u * p < osplit) {                                     // the pair straddles the split
                        const uint32_t i1 = LUT ? __ldg(a.lut + 2u * p + 1u) : (2u * p + 1u) * R;
                        return pack_s16x2_sat((int)ld_l2(st + 2u * p), enc_s32(i1 < n_in ? fr[i1 + 4u * __umulhi(i1, mq)] : qnan));
                    }
                    float x0, x1;
                    if (LUT) {
                        const uint2 ii = __ldg(lut2 + p);
                        x0 = ii.x < n_in ? fr[ii.x + 4u * __umulhi(ii.x, mq)] : qnan;
                        x1 = ii.y < n_in ? fr[ii.y + 4u * __umulhi(ii.y, mq)] : qnan;
                    } else {
                        const uint32_t i0 = 2u * p * R, i1 = i0 + R;
                        x0 = fr[i0 + 4u * __umulhi(i0, mq)];
                        x1 = fr[i1 + 4u * __umulhi(i1, mq)];
                    }
                    return pack_s16x2_sat(enc_s32(x0), enc_s32(x1));
                };
                // two pairs per trip, by hand: `#pragma unroll 2` made the compiler divide by blockDim.x for the trip count
                for (uint32_t p = threadIdx.x; p < nfull; p += 2u * bd) {
                    const uint32_t pb = p + bd;
                    const uint32_t va = pair(p), vb = pb < nfull ? pair(pb) : 0u;
                    st_32(dstf + 2u * p, va, 1);                               // streaming stores: frames are written once
                    if (pb < nfull) st_32(dstf + 2u * pb, vb, 1);
                }
                if ((n_out & 1u) && threadIdx.x == 0u) {                       // the odd last output
                    const uint32_t o = n_out - 1u, i = LUT ? __ldg(a.lut + o) : o * R;
                    dstf[o] = o < osplit ? ld_l2(st + o) : enc_s16(i < n_in ? fr[i + 4u * __umulhi(i, mq)] : qnan);
                }
            } else {                                                           // unaligned caller buffer: 16-bit stores
                for (uint32_t o = threadIdx.x; o < n_out; o += blockDim.x) {
                    const uint32_t i = LUT ? __ldg(a.lut + o) : o * g.stride;
                    dstf[o] = o < osplit ? ld_l2(st + o) : enc_s16(i < n_in ? fr[i + 4u * __umulhi(i, a.magic_q)] : qnan);
                }
            }
        }

        if (METER) {
            // lane <-> slot; the quanta carried over from earlier pushes were metered when they were appended
            const uint32_t m_warps = (T + 31u) >> 5;
            const uint32_t rot = m_warps < nw ? (it * m_warps) % nw : 0u;
            for (uint32_t sl = (((warp + nw - rot) % nw) << 5) + lane; sl < T; sl += blockDim.x) {
                uint32_t k = 0;
                while (k + 1u < nf && sl >= m.qstart[k + 1u]) k++;
                const uint32_t j = sl - m.qstart[k];
                if (j < m.fd[k].ring_count) continue;
                a.rms_at[m.fd[k].pos + (j - m.fd[k].ring_count)] = quantum_rms_cvt(reinterpret_cast<const float4 *>(arena + (size_t)sl * pitch), Q);
            }
        }
        if (!more) break;
        __syncthreads();                                   // the arena and this step's meta buffer are free again
        f0 = f_next;
    }
}

// ---- 6 (usual form): convert, one frame per step, descriptors in registers -----------------------------------------
// The common geometry -- one ~60 ms frame per 128-thread step, at most 64 quanta per frame -- needs none of the
// shared-memory bookkeeping of k_agg_convert_pipe: ncu counted ~600 of its 1 570 warp-instructions per frame in
// per-step address arithmetic for the meta buffers and ~100 in building them.  Here every thread holds the frame's
// descriptor in registers, lane l of every warp holds the push index of the frame's l-th (and l+32-nd) push quantum
// (turned into source addresses at the top of a trip and broadcast with a shuffle when the warp copies that slot);
// descriptors are fetched two frames ahead and index lists one frame ahead, while copies are in flight.  No meta
// buffers, two block barriers per frame, a persistent grid (the CTAs resident at once).
template <bool METER, bool LUT>
__global__ void __launch_bounds__(128, 10)
k_agg_convert_one(AggGeom g, AggConvertArgs a)
{
    pdl_enter();
    const CtlHead head = ctl_head(a.ctl);
    if (head.status != 0u) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float *arena = reinterpret_cast<float *>(smem_raw);
    const uint32_t Q = g.quantum, pitch = Q + 4u, gpq = Q >> 2, mq = a.magic_q, bd = blockDim.x;
    const uint32_t n_frames = head.n_frames;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nw = bd >> 5;
    uint32_t f = blockIdx.x;
    if (f >= n_frames) return;

    // What a thread keeps of a frame: its descriptor, and in lane l the push index of the frame's l-th and (l+32)-nd
    // push quantum (slots ring_count + l, ring_count + l + 32).  The descriptor is fetched TWO frames ahead and the
    // index list one frame ahead, so that the list's address (it depends on the descriptor) is known when the list is
    // asked for, and nothing is computed from a prefetched value before the next trip: a thread never waits for a
    // prefetch in the middle of a trip (ncu: a third of all stall samples sat on that dependency before).
    struct Desc {
        unsigned long long off;
        uint32_t session, nq, rc, pos;
    };
    struct Frame {
        Desc d;
        uint32_t idx0, idx1, n_out, osplit;                                   // outputs below osplit were converted on arrival
    };
    auto fetch_desc = [&](uint32_t fi) -> Desc {
        const uint4 raw = __ldcg(reinterpret_cast<const uint4 *>(a.fdesc + fi));   // {session, nq, ring_count, pos}
        return Desc{ld_l2(a.fdesc_off + fi), raw.x, raw.y, raw.z, raw.w};
    };
    auto fetch_list = [&](const Desc &d) -> Frame {
        const uint32_t npush = d.nq - d.rc;
        Frame fr;
        fr.d = d;
        fr.idx0 = lane < npush ? ld_l2(a.sorted + d.pos + lane) : 0u;
        fr.idx1 = lane + 32u < npush ? ld_l2(a.sorted + d.pos + lane + 32u) : 0u;
        fr.n_out = __ldg(a.nout_of + d.nq);
        fr.osplit = d.rc ? __ldg(a.out_lo + d.rc) : 0u;
        return fr;
    };
    const uint32_t stride_f = gridDim.x;
    __shared__ const float *ptab[4][64];                                       // per warp: where each slot's quantum comes from
    Frame cur = fetch_list(fetch_desc(f));
    Desc nxt_d = cur.d;                                                        // descriptor of the frame after `cur`
    if (f + stride_f < n_frames) nxt_d = fetch_desc(f + stride_f);
    const uint32_t dst_lane = smem_u32(arena) + 16u * lane, pitch_b = pitch * 4u;
    for (uint32_t it = 0;; it++) {
        const uint32_t nq = cur.d.nq, rc = cur.d.rc;
        // lane l: where slots l and l + 32 come from (the first ring_count slots were carried over from earlier
        // pushes; the rest are this push's quanta), already offset to this lane's 16 bytes
        const uint32_t j0 = lane - rc, j1 = lane + 32u - rc;                 // push quantum behind slot lane / lane + 32
        const uint32_t a0 = __shfl_sync(0xFFFFFFFFu, cur.idx0, j0 & 31u);
        const uint32_t b0 = __shfl_sync(0xFFFFFFFFu, cur.idx0, j1 & 31u), b1 = __shfl_sync(0xFFFFFFFFu, cur.idx1, j1 & 31u);
        // (slots below rc were converted when they arrived -- a.staged -- and are not fetched at all)
        const float *src0 = a.samples + (size_t)a0 * Q;
        const float *src1 = a.samples + (size_t)(j1 < 32u ? b0 : b1) * Q;
        // ---- all quanta of the frame in flight at once: one warp-wide 16-byte-per-lane copy per 512-byte quantum.
        // Four slots per trip, addresses first and the four copies back to back (the compiler pads every isolated
        // LDGSTS with three dummy instructions; 32 instructions per quantum in the one-at-a-time form, per ncu).
        // Every warp parks its copy of the 64 source addresses in its own corner of shared memory (no block barrier:
        // the table is private to the warp) and reads them back with warp-uniform, i.e. broadcast, loads: shuffling
        // 64-bit addresses cost 20 instructions per quantum (two shuffles, two selects, a divergence guard ...), ncu.
        ptab[warp][lane] = src0;
        ptab[warp][lane + 32u] = src1;
        __syncwarp();
        if (gpq == 32u) {
            for (uint32_t s0 = rc + warp; s0 < nq; s0 += 4u * nw) {
                const float *src[4];
#pragma unroll
                for (int u = 0; u < 4; u++) src[u] = ptab[warp][min(s0 + (uint32_t)u * nw, 63u)] + 4u * lane;
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const uint32_t sl = s0 + (uint32_t)u * nw;
                    if (sl < nq) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_lane + sl * pitch_b), "l"(src[u]) : "memory");
                }
            }
        } else {
            for (uint32_t sl = rc + warp; sl < nq; sl += nw) {
                const float *src = ptab[warp][sl];
#pragma unroll 1
                for (uint32_t gi = lane; gi < gpq; gi += 32u) cp_async_16(arena + (size_t)sl * pitch + 4u * gi, src + 4u * gi);
            }
        }
        __syncwarp();                                      // the next trip rewrites the table
        // The outputs of the carried-over quanta (converted when they arrived, a.staged) ride in the SAME round trip: the
        // arena slots below rc hold nothing, and rc quanta's outputs fit there at any ratio >= quantum / (quantum + 4) / 2
        // (8 kHz and up).  Fetching them in `picks`, after the wait, cost every such frame a second trip to DRAM -- the
        // whole time of a streaming flush (1 new quantum, 22 carried over: 1.5 ms per 2^20 frames of 4.4 KB traffic each).
        const uint32_t osplit = cur.osplit, pre_chunks = (osplit * 2u + 15u) >> 4;
        const bool pre_smem = pre_chunks * 16u <= rc * pitch_b;                 // block-uniform
        const int16_t *st = a.staged + (size_t)cur.d.session * a.nout_pad;
        if (pre_smem)
            for (uint32_t c = threadIdx.x; c < pre_chunks; c += bd)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(arena) + 16u * c), "l"(st + 8u * c) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
        // the frame after this one: its descriptor is here already (fetched a trip ago), so its index list can be asked
        // for now; and the descriptor of the frame after that.  Both land while this frame's copies are in flight.
        const uint32_t f_next = f + stride_f;
        const bool more = f_next < n_frames;                                   // block-uniform
        Frame nxt = cur;
        Desc ahead = nxt_d;
        if (more) {
            nxt = fetch_list(nxt_d);
            if (f_next + stride_f < n_frames) ahead = fetch_desc(f_next + stride_f);
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();

        // ---- pick + encode + store (see k_agg_convert_pipe); pointers and indices advance by addition.  `tid` of `nthr`
        // threads share the frame's pairs: all of the CTA, or -- with the meter -- the warps that are not metering.
        auto picks = [&](uint32_t tid, uint32_t nthr) {
            const uint32_t n_in = nq * Q, n_out = cur.n_out;
            int16_t *dstf = a.out + cur.d.off;
            const float qnan = __int_as_float(0x7FC00000);
            // outputs below `osplit` belong to quanta that were converted when they arrived: they sit at the front of
            // the arena (or, at ratios below 1/2, still in a.staged)
            const uint32_t *pre32 = reinterpret_cast<const uint32_t *>(arena);
            const int16_t *pre16 = reinterpret_cast<const int16_t *>(arena);
            auto staged16 = [&](uint32_t o) -> int16_t { return pre_smem ? pre16[o] : ld_l2(st + o); };
            if (a.out_aligned16) {
                const uint32_t nfull = n_out >> 1, R = g.stride;
                // the whole 16-byte pieces of that prefix are a straight copy (a streaming flush is 22 carried-over quanta
                // and one new one: through the pair loop below the copy cost 30 instructions per pair, and the kernel was
                // issue-bound at 1 090 warp-instructions per frame, ncu)
                const uint32_t p0 = pre_smem ? (osplit >> 3) << 2 : 0u;         // first pair the pair loop handles
                for (uint32_t c = tid; c < (p0 >> 2); c += nthr) {
                    const uint4 v = reinterpret_cast<const uint4 *>(arena)[c];
                    asm volatile("st.global.cs.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(reinterpret_cast<uint4 *>(dstf) + c), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
                }
                // the few pairs left of the prefix (less than one 16-byte piece, and the pair that straddles the split if
                // osplit is odd) each by one thread; kept out of the main loop, whose pairs then need no test
                const uint32_t psplit = min((osplit + 1u) >> 1, nfull);
                for (uint32_t p = p0 + tid; p < psplit; p += nthr) {
                    uint32_t v;
                    if (2u * p + 1u < osplit) {
                        v = pre_smem ? pre32[p] : ld_l2(reinterpret_cast<const uint32_t *>(st) + p);
                    } else {
                        const uint32_t i1 = LUT ? __ldg(a.lut + 2u * p + 1u) : (2u * p + 1u) * R;
                        v = pack_s16x2_sat((int)staged16(2u * p), enc_s32(i1 < n_in ? arena[i1 + 4u * __umulhi(i1, mq)] : qnan));
                    }
                    reinterpret_cast<uint32_t *>(dstf)[p] = v;
                }
                uint32_t *dp = reinterpret_cast<uint32_t *>(dstf) + psplit + tid;
                const uint2 *lp = reinterpret_cast<const uint2 *>(a.lut) + psplit + tid;
                uint32_t i0 = 2u * (psplit + tid) * R;
                const uint32_t di = 2u * nthr * R;                              // index step from pair p to pair p + nthr
                auto pair = [&](uint32_t ia, const uint2 *l) -> uint32_t {
                    float x0, x1;
                    if (LUT) {
                        const uint2 ii = __ldg(l);
                        x0 = ii.x < n_in ? arena[ii.x + 4u * __umulhi(ii.x, mq)] : qnan;
                        x1 = ii.y < n_in ? arena[ii.y + 4u * __umulhi(ii.y, mq)] : qnan;
                    } else {
                        const uint32_t ib = ia + R;
                        x0 = arena[ia + 4u * __umulhi(ia, mq)];
                        x1 = arena[ib + 4u * __umulhi(ib, mq)];
                    }
                    return pack_s16x2_sat(enc_s32(x0), enc_s32(x1));
                };
                for (uint32_t p = psplit + tid; p < nfull; p += 2u * nthr, dp += 2u * nthr, lp += 2u * nthr, i0 += 2u * di) {
                    const bool second = p + nthr < nfull;
                    const uint32_t va = pair(i0, lp), vb = second ? pair(i0 + di, lp + nthr) : 0u;
                    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(dp), "r"(va) : "memory");   // streaming: frames are written once
                    if (second) asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(dp + nthr), "r"(vb) : "memory");
                }
                if ((n_out & 1u) && tid == 0u) {
                    const uint32_t o = n_out - 1u, i = LUT ? __ldg(a.lut + o) : o * R;
                    dstf[o] = o < osplit ? staged16(o) : enc_s16(i < n_in ? arena[i + 4u * __umulhi(i, mq)] : qnan);
                }
            } else {
                for (uint32_t o = tid; o < n_out; o += nthr) {
                    const uint32_t i = LUT ? __ldg(a.lut + o) : o * g.stride;
                    dstf[o] = o < osplit ? staged16(o) : enc_s16(i < n_in ? arena[i + 4u * __umulhi(i, mq)] : qnan);
                }
            }
        };
        if (!METER) {
            picks(threadIdx.x, bd);
        } else {
            // lane <-> slot; carried-over quanta were metered when they were appended.  A quantum's sum of squares is a
            // 128-long dependent chain (~1 300 clocks) during which the arena cannot be refilled, so the metering warps
            // start on it at once and the OTHER warps share the picks; which warps meter rotates from frame to frame so
            // that the chains do not always land on the same scheduler.
            const uint32_t m_warps = (nq + 31u) >> 5;
            auto meter = [&](uint32_t sl) {
                if (sl < nq && sl >= rc) a.rms_at[cur.d.pos + (sl - rc)] = quantum_rms_cvt(reinterpret_cast<const float4 *>(arena + (size_t)sl * pitch), Q);
            };
            if (m_warps < nw) {
                const uint32_t wrel = (warp + nw - (it * m_warps) % nw) % nw;
                if (wrel < m_warps) meter((wrel << 5) + lane);
                else picks(((wrel - m_warps) << 5) + lane, (nw - m_warps) << 5);
            } else {
                picks(threadIdx.x, bd);
                for (uint32_t sl = threadIdx.x; sl < nq; sl += bd) meter(sl);
            }
        }
        if (!more) break;
        __syncthreads();                                   // the arena is free again
        f = f_next;
        cur = nxt;
        nxt_d = ahead;
    }
}

// ---- 6 (streaming form): convert, one frame per WARP ---------------------------------------------------------------
// A server that forwards quanta as they arrive pushes ONE quantum per session at a time: a flushing frame is then
// ~22 carried-over quanta -- 1.9 KB of outputs that were converted on arrival and only need copying -- and one new
// quantum.  k_agg_convert_one gives such a frame a whole CTA and an arena sized for the longest frame: ten frames in
// flight per SM at 2.4 KB each, 1.39 ms per 2^20 frames (0.52 of the HBM peak, bound by bytes in flight; ncu).  Here a
// warp owns a frame: the carried-over outputs go global -> registers -> global in 16-byte pieces, the new quanta
// pass through a four-slot window of shared memory private to the warp (outputs [out_lo[j], out_lo[j + 4]) need
// quanta j .. j + 3 only), no block barrier anywhere, 40 frames in flight per SM.  Any frame is handled correctly --
// a long one takes several windows, one after the other -- so the host may pick this kernel on a guess (the previous
// push had short lists) without anything depending on the guess being right.
constexpr uint32_t kWarpWindow = 4;
template <bool METER, bool LUT>
__global__ void __launch_bounds__(128, METER ? 8 : 10)
k_agg_convert_warp(AggGeom g, AggConvertArgs a)
{
    pdl_enter();
    const CtlHead head = ctl_head(a.ctl);
    if (head.status != 0u) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t W = kWarpWindow;
    const uint32_t Q = g.quantum, pitch = Q + 4u, gpq = Q >> 2, mq = a.magic_q;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nw = blockDim.x >> 5;
    float *win = reinterpret_cast<float *>(smem_raw) + (size_t)warp * W * pitch;
    const uint32_t n_frames = head.n_frames;
    const float qnan = __int_as_float(0x7FC00000);
    for (uint32_t f = blockIdx.x * nw + warp; f < n_frames; f += gridDim.x * nw) {
        const AggFrameDesc d = ld_l2(a.fdesc + f);                             // {session, nq, ring_count, pos}
        const unsigned long long off = ld_l2(a.fdesc_off + f);
        const uint32_t nq = d.nq, rc = d.ring_count, n_in = nq * Q;
        const uint32_t n_out = __ldg(a.nout_of + nq), osplit = rc ? __ldg(a.out_lo + rc) : 0u;
        int16_t *dst = a.out + off;
        const int16_t *st = a.staged + (size_t)d.session * a.nout_pad;
        // the first new quanta are asked for before anything else is waited for
        const uint32_t cnt0 = min(W, nq - rc);
        uint32_t idx = lane < cnt0 ? ld_l2(a.sorted + d.pos + lane) : 0u;
        // ---- the carried-over outputs [0, osplit): a straight copy, four 16-byte pieces per lane per trip, loads first
        uint32_t done = 0;
        if (a.out_aligned16) {
            const uint32_t chunks = osplit >> 3;
            const uint4 *src4 = reinterpret_cast<const uint4 *>(st);
            uint4 *dst4 = reinterpret_cast<uint4 *>(dst);
            for (uint32_t c0 = lane; c0 < chunks; c0 += 128u) {
                uint4 v[4];
#pragma unroll
                for (uint32_t u = 0; u < 4u; u++) if (c0 + 32u * u < chunks) v[u] = __ldcg(src4 + c0 + 32u * u);
#pragma unroll
                for (uint32_t u = 0; u < 4u; u++)
                    if (c0 + 32u * u < chunks)
                        asm volatile("st.global.cs.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(dst4 + c0 + 32u * u), "r"(v[u].x), "r"(v[u].y), "r"(v[u].z), "r"(v[u].w) : "memory");
            }
            done = chunks << 3;
        }
        for (uint32_t o = done + lane; o < osplit; o += 32u) dst[o] = ld_l2(st + o);
        // ---- the new quanta, W at a time through the warp's window
        for (uint32_t j0 = rc; j0 < nq; j0 += W) {
            const uint32_t cnt = min(W, nq - j0);
            if (j0 != rc) idx = lane < cnt ? ld_l2(a.sorted + d.pos + (j0 - rc) + lane) : 0u;
            __syncwarp();                                                      // the previous window's picks are over
            for (uint32_t t = 0; t < cnt; t++) {
                const float *src = a.samples + (size_t)__shfl_sync(0xFFFFFFFFu, idx, t) * Q;
                for (uint32_t gi = lane; gi < gpq; gi += 32u) cp_async_16(win + (size_t)t * pitch + 4u * gi, src + 4u * gi);
            }
            cp_async_wait_all();
            __syncwarp();
            // outputs whose pick falls in these quanta; the last window also takes the outputs whose pick lies beyond
            // the frame (undefined in the reference, i.e. NaN, i.e. 0)
            const uint32_t o_lo = __ldg(a.out_lo + j0), o_hi = j0 + cnt < nq ? __ldg(a.out_lo + j0 + cnt) : n_out;
            const uint32_t base = j0 * Q;
            for (uint32_t o = o_lo + lane; o < o_hi; o += 32u) {
                const uint32_t is = LUT ? __ldg(a.lut + o) : o * g.stride;
                float x = qnan;
                if (is < n_in) {
                    const uint32_t li = is - base;
                    x = win[li + 4u * __umulhi(li, mq)];
                }
                dst[o] = enc_s16(x);
            }
            if (METER && lane < cnt)
                a.rms_at[d.pos + (j0 - rc) + lane] = quantum_rms_cvt(reinterpret_cast<const float4 *>(win + (size_t)lane * pitch), Q);
        }
    }
}

// ---- 7: levels ---------------------------------------------------------------------------------------------------
// App.tsx:271 along every touched session's FLUSHED quanta, in arrival order: first the quanta that were
// waiting from earlier pushes (their rms was stored in ring_rms when they were appended), then this push's.
__global__ void __launch_bounds__(128)
k_agg_levels(AggGeom g, const uint32_t *__restrict__ touched, const uint32_t *__restrict__ count, const uint32_t *__restrict__ cursor,
             const uint32_t *__restrict__ fill, const uint32_t *__restrict__ flushed, const double *__restrict__ ring_rms,
             const double *__restrict__ rms_at, double *__restrict__ level, const AggCtl *__restrict__ ctl)
{
    pdl_enter();
    const CtlHead head = ctl_head(ctl);
    if (head.status != 0u) return;
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= head.n_touched) return;
    const uint32_t nfl = ld_l2(flushed + k);
    if (nfl == 0u) return;
    const uint32_t s = ld_l2(touched + k), base = ld_l2(cursor + s) - ld_l2(count + s), c0 = ld_l2(fill + s);
    double lv = ld_l2(level + s);
    bool poisoned = lv != lv;
    auto step = [&](double rms) {
        poisoned |= rms != rms;
        const double dec = __dmul_rn(lv, 0.85);
        lv = rms > dec ? rms : dec;
        if (poisoned) lv = __longlong_as_double(0x7FF8000000000000ll);
    };
    // the chain is sequential, the loads are not: fetch 8 rms values per round trip
    auto run = [&](const double *src, uint32_t cnt) {
        for (uint32_t j0 = 0; j0 < cnt; j0 += 8u) {
            double r[8];
#pragma unroll
            for (uint32_t u = 0; u < 8u; u++) r[u] = j0 + u < cnt ? ld_l2(src + j0 + u) : 0.0;
#pragma unroll
            for (uint32_t u = 0; u < 8u; u++) if (j0 + u < cnt) step(r[u]);
        }
    };
    run(ring_rms + (size_t)s * g.half_quanta, c0);
    run(rms_at + base, nfl);
    level[s] = lv;
}

// ---- 8: append ------------------------------------------------------------------------------------------------------
// Quanta that did not flush are converted where they stand (AggAppendArgs below).  A CTA brings TILE (32 or 64)
// consecutive quanta of the push into shared memory with 16-byte async copies (slot pitch quantum + 4 floats) --
// picking straight from global memory was latency-bound, 300 vs 230 us per 2^20 quanta -- and its warps then
// pick + encode each waiting quantum's outputs into staged[session][].  With the meter on, lane t first meters
// slot t exactly like kernel 2c: the rms must be accumulated in sample order by ONE lane (handing a running sum
// from lane to lane costs ~15 warp-instructions per sample).
// What the commit step needs (it rides in the append kernel: one launch fewer per push).
struct AggCommitArgs {
    const uint32_t *touched;
    uint32_t *count, *fill;
    double *last_send;
    const uint32_t *new_fill;
    const double *new_last;
    AggCtl *ctl_next;                // the NEXT push's control block: cleared here
};

// Where a quantum that does not flush goes.  A session's waiting quanta are not kept as samples: quantum j of a
// frame owns the outputs [out_lo[j], out_lo[j + 1]) whatever the frame's final length (the pick sequence does not
// depend on it, App.tsx:26-30), so they are picked and encoded on arrival and only those ~43 int16 (of 128 floats at
// 48 kHz) are kept: `staged[session][output]`.  A flush then copies that prefix instead of converting 16 KB of floats.
struct AggAppendArgs {
    int16_t *staged;                 // [max_sessions][nout_pad]
    const uint32_t *out_lo;          // [half_quanta + 1]
    const uint32_t *lut;             // pick table (one entry of padding), or null for an integer ratio
    uint32_t nout_pad, half_quanta, stride;
    uint32_t pairs, magic_pairs;     // aligned output PAIRS a quantum's outputs can touch (the most over all positions), >= 2
};

// Where one waiting quantum goes: session row (0xFFFFFFFF: the quantum flushed, nothing to do), first sample index of
// its quantum in the frame, its outputs [olo, ohi).
struct AggStageDesc { uint32_t row, i0, olo, ohi; };
__device__ __forceinline__ AggStageDesc agg_stage_desc(const AggAppendArgs &ap, uint32_t slot, uint32_t quantum)
{
    const uint32_t s = slot / ap.half_quanta, j = slot - s * ap.half_quanta;
    return AggStageDesc{s, j * quantum, __ldg(ap.out_lo + j), __ldg(ap.out_lo + j + 1u)};
}

// TILE quanta per block step, TILE * 4 threads (64 / 256 for large pushes; 32 / 128 keeps a small push in one wave).
// The picks: a work item is one ALIGNED PAIR of outputs of one quantum of the tile, items dealt to threads round-robin
// (item f -> quantum f / pairs, pair f % pairs).  The warp-per-quantum form this replaces ran two passes of 32 lanes
// over ~43 outputs with 16-bit stores and ~26 instructions of per-quantum preamble per warp: 82 of the kernel's 109
// warp-instructions per quantum (ncu, 2^20 quanta that all wait: issue-active 50 %, 211 us).  Now 43 of 72, 191 us.
// What is left is not the kernel's to give: a software-pipelined persistent form (slots read two tiles ahead, tile
// k+1 in flight into a second buffer while tile k is picked) measured 206 us, the plain ring copy this design
// replaced 176 us for 1.7x the bytes -- every form lands on ~2^20 scattered row activations per push (one 86-byte
// write per session, plus the read that completes its partial sectors: DRAM moves 0.63 + 0.15 GB for 0.54 + 0.09).
// SUB tiles per trip (1 or 4): the slots of SUB * TILE quanta are read in one coalesced request and a trip whose quanta
// all flushed costs one round trip, not SUB -- a push in which every quantum flushes (65 536 sessions x 23) is nothing
// but such trips, 24 us of them with SUB = 1.
// (register budget stated: left alone, ptxas settles on 32 registers and spills the prefetched commit item)
template <bool METER, uint32_t TILE, uint32_t SUB>
__global__ void __launch_bounds__(TILE * 4, (METER ? 256 : 384) / TILE)
k_agg_append(const float *__restrict__ samples, const uint32_t *__restrict__ dst_slot, AggAppendArgs ap,
             double *__restrict__ ring_rms, uint32_t n, uint32_t quantum, const AggCtl *__restrict__ ctl, AggCommitArgs cm)
{
    pdl_enter();
    // ---- 9: commit.  fill / lastSend of every touched session take their new values (only if the push stands) and
    // the per-push counters are cleared (always: the next push starts from clean counters).  The append work neither
    // reads nor writes any of this, so where the commit sits in the kernel is only a matter of overlap: the thread's
    // first commit item is fetched now and stored after the append work, so that its memory round trips run under the
    // append's; items beyond the grid's size are done in a loop at the end.  (One call site: as a lambda called from
    // two places the commit was not inlined, its captures went through the stack, and the very first thing a thread
    // did was to wait for the prefetch it was supposed to overlap -- 5 % of the kernel's stall samples.)
    const CtlHead head = ctl_head(ctl);
    const bool ok = head.status == 0u;
    const uint32_t nt = head.n_touched, k0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (blockIdx.x == 0 && threadIdx.x == 0) *cm.ctl_next = AggCtl{};
    uint32_t c_s = 0, c_fill = 0;
    double c_last = 0.0;
    if (k0 < nt) {
        c_s = ld_l2(cm.touched + k0);
        if (ok) { c_fill = ld_l2(cm.new_fill + k0); c_last = ld_l2(cm.new_last + k0); }
    }
    if (ok) {
        constexpr uint32_t DEAD = 0xFFFFFFFFu, NW = TILE / 8u;     // warps per block
        const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
        const uint32_t granules = quantum >> 2;
        extern __shared__ __align__(128) unsigned char smem_raw[];
        const uint32_t pitch = quantum + 4u;
        float *arena = reinterpret_cast<float *>(smem_raw);
        static_assert(SUB >= 1 && SUB <= 4, "one slot per thread");
        constexpr uint32_t ST = TILE * SUB;                        // quanta per trip
        __shared__ uint32_t slot_s[ST];
        __shared__ AggStageDesc desc_s[ST];                        // worked out once per quantum (one thread each), not once per item
        for (uint32_t s0 = blockIdx.x * ST; s0 < n; s0 += gridDim.x * ST) {
            uint32_t slot = DEAD;
            if (threadIdx.x < ST && s0 + threadIdx.x < n) slot = ld_l2(dst_slot + s0 + threadIdx.x);
            // (a barrier: the previous trip's picks are over, its shared memory may be rewritten)
            if (__syncthreads_count(slot != DEAD) == 0) continue;
            if (threadIdx.x < ST) {
                slot_s[threadIdx.x] = slot;
                desc_s[threadIdx.x] = slot != DEAD ? agg_stage_desc(ap, slot, quantum) : AggStageDesc{DEAD, 0u, 0u, 0u};
            }
            for (uint32_t sub = 0; sub < SUB; sub++) {
                const uint32_t q0 = sub * TILE, i0 = s0 + q0;
                // (a barrier again: descriptors visible, the previous tile's picks over) does any quantum of this tile wait?
                if (__syncthreads_or(threadIdx.x / TILE == sub && slot != DEAD) == 0) continue;
                if (granules == 32u) {                   // the 128-sample quantum: one warp-wide copy per quantum, no division
                    for (uint32_t t = warp; t < TILE; t += NW)
                        if (desc_s[q0 + t].row != DEAD) cp_async_16(arena + (size_t)t * pitch + 4u * lane, samples + (size_t)(i0 + t) * quantum + 4u * lane);
                } else {
                    for (uint32_t idx = threadIdx.x; idx < TILE * granules; idx += blockDim.x) {
                        const uint32_t t = idx / granules, gi = idx - t * granules;
                        if (desc_s[q0 + t].row != DEAD) cp_async_16(arena + (size_t)t * pitch + 4u * gi, samples + (size_t)(i0 + t) * quantum + 4u * gi);
                    }
                }
                cp_async_wait_all();
                __syncthreads();
                auto picks = [&](uint32_t tid, uint32_t nthr) {
                    for (uint32_t f = tid; f < TILE * ap.pairs; f += nthr) {
                        const uint32_t t = __umulhi(f, ap.magic_pairs), q = f - t * ap.pairs;
                        const AggStageDesc d = desc_s[q0 + t];
                        const uint32_t p = (d.olo & ~1u) + 2u * q;              // the pair (p, p + 1), p even
                        if (d.row == DEAD || p >= d.ohi) continue;
                        const bool lo_ok = p >= d.olo, hi_ok = p + 1u < d.ohi;  // at least one holds
                        uint32_t ia, ib;
                        if (ap.lut) {
                            const uint2 ii = __ldg(reinterpret_cast<const uint2 *>(ap.lut + p));
                            ia = ii.x; ib = ii.y;
                        } else {
                            ia = p * ap.stride; ib = ia + ap.stride;
                        }
                        // picks of [olo, ohi) lie inside the quantum by the definition of out_lo
                        const float *src = arena + (size_t)t * pitch;
                        const float x0 = lo_ok ? src[ia - d.i0] : 0.0f, x1 = hi_ok ? src[ib - d.i0] : 0.0f;
                        const uint32_t v = pack_s16x2_sat(enc_s32(x0), enc_s32(x1));
                        int16_t *dst = ap.staged + (size_t)d.row * ap.nout_pad + p;
                        if (lo_ok && hi_ok) *reinterpret_cast<uint32_t *>(dst) = v;
                        else if (lo_ok) dst[0] = (int16_t)(v & 0xFFFFu);
                        else dst[1] = (int16_t)(v >> 16);
                    }
                };
                if (METER) {
                    // threads 0 .. TILE-1 meter slot t (a 128-long dependent chain each); the other three quarters of
                    // the block share the picks meanwhile
                    if (threadIdx.x < TILE) {
                        const uint32_t sl = slot_s[q0 + threadIdx.x];
                        if (sl != DEAD)
                            ring_rms[sl] = quantum_rms_cvt(reinterpret_cast<const float4 *>(arena + (size_t)threadIdx.x * pitch), quantum);
                    } else {
                        picks(threadIdx.x - TILE, blockDim.x - TILE);
                    }
                } else {
                    picks(threadIdx.x, blockDim.x);
                }
            }
        }
    }
    if (k0 < nt) {
        cm.count[c_s] = 0u;
        if (ok) { cm.fill[c_s] = c_fill; cm.last_send[c_s] = c_last; }
    }
    for (uint32_t k = k0 + gridDim.x * blockDim.x; k < nt; k += gridDim.x * blockDim.x) {
        const uint32_t s = ld_l2(cm.touched + k);
        cm.count[s] = 0u;
        if (ok) { cm.fill[s] = ld_l2(cm.new_fill + k); cm.last_send[s] = ld_l2(cm.new_last + k); }
    }
}

__global__ void k_agg_reset(uint32_t *__restrict__ fill, double *__restrict__ last_send, uint32_t first, uint32_t n, double now_ms)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) { fill[first + k] = 0u; last_send[first + k] = now_ms; }
}

}  // namespace pcmb200
