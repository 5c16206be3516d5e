// pcm_aggregator.inl -- host side of the device-side capture aggregator (included by pcm_b200.cu).
//
// Restates the reference's aggregator (App.tsx:148-157, :273-288) for many sessions: per session,
// 128-sample quanta are appended to a buffer; the quantum that arrives with now - lastSend >= desiredMs
// flushes the WHOLE buffer (a multiple of 128 samples, e.g. 23 quanta = 2944 samples -> 981 outputs at
// 48 kHz) through floatTo16BitPCM(resampleTo16k(buf, rate)) and resets buffer and lastSend.  The wall
// clock is injected: every quantum carries its arrival time.
//
// All per-session state (fill, lastSend, level, the buffer itself) lives in HBM, and the flush plan of
// a push is computed by kernels (agg_kernels.cuh): the host enqueues a fixed sequence of launches and
// reads back {status, n_frames, out_total}.  A rejected push changes nothing (the plan is scratch until
// the commit kernel runs).
#include "agg_kernels.cuh"

struct pcmb200_aggregator {
    AggGeom geo{};
    uint32_t max_push = 0, nout_max = 0;
    size_t out_cap = 0;
    double rate = 0;
    bool meter = false;
    // persistent device state
    int16_t *staged = nullptr;             // [max_sessions][nout_pad]: the waiting quanta of a session, ALREADY converted
    uint32_t *out_lo = nullptr;            // [half_quanta + 1] first output whose pick falls at or after quantum j
    uint32_t nout_pad = 0;                 // nout_max rounded up to 8 samples
    double *ring_rms = nullptr;            // [max_sessions][half_quanta] rms of every waiting quantum (meter)
    uint32_t *fill = nullptr, *count = nullptr, *cursor = nullptr;   // [max_sessions]
    double *last_send = nullptr, *level = nullptr;                   // [max_sessions]
    uint32_t *nout_of = nullptr;           // [half_quanta + 1] output length of a frame of nq quanta
    uint32_t *lut = nullptr;               // one pick table for every frame length (see create)
    // per-push scratch, sized for max_push quanta
    uint32_t *d_sessions = nullptr;        // staging when the caller's arrays are host memory
    double *d_arrival = nullptr;
    float *d_stage = nullptr;
    uint32_t *touched_idx = nullptr;       // [touched] the push index of the session's first quantum
    uint32_t *touched = nullptr, *sorted = nullptr, *fl_nq = nullptr, *fl_ring = nullptr, *fl_pos = nullptr,
             *dst_slot = nullptr, *new_fill = nullptr, *flushed = nullptr;
    double *new_last = nullptr, *rms_at = nullptr;
    uint32_t *sums_a = nullptr, *sums_b = nullptr;
    AggFrameDesc *fdesc = nullptr;
    unsigned long long *fdesc_off = nullptr;
    uint32_t *f_session = nullptr, *f_len = nullptr;     // frame arrays when the caller's are host memory
    unsigned long long *f_off = nullptr;
    int16_t *d_out = nullptr;              // [out_cap] when the caller's output is host memory
    AggCtl *ctl = nullptr;                 // device: two blocks, used by alternate pushes (see agg_enqueue)
    int ctl_idx = 0;
    AggCtl *h_ctl = nullptr;               // pinned copy
    unsigned warp_grid[2] = {0, 0};        // k_agg_convert_warp: resident CTAs on the whole device, without / with the meter
    size_t convert_warp_smem = 0;
    uint32_t pairs = 2;                    // aligned output pairs per quantum, at most (append kernel)
    uint32_t prev_n = 0;                   // quanta in the previous push (with h_ctl->n_touched: which walk kernel to take)
    // convert kernel geometry
    uint32_t G = 1;
    size_t convert_pipe_smem = 0, convert_one_smem = 0, append_smem = 0;
    unsigned one_grid = 0;                 // k_agg_convert_one: resident CTAs on the whole device for ...
    int one_grid_key = -1;                 // ... this (threads per CTA, meter) combination
    bool pending = false;                  // a push was enqueued and its result not yet collected
};

namespace {

int agg_status_to_error(pcmb200_ctx *ctx, uint32_t st)
{
    if (st == 0) return PCMB200_OK;
    if (st & AGG_ST_BAD_SESSION) return fail(ctx, PCMB200_ERR_INVALID, "pcmb200_aggregator_push: session id out of range (push rejected, nothing changed)");
    if (st & AGG_ST_FRAME_TOO_LONG)
        return fail(ctx, PCMB200_ERR_CAPACITY, "pcmb200_aggregator_push: a session's buffer exceeds max_frame_quanta (no flush arrived); push rejected, nothing changed");
    if (st & AGG_ST_TOO_MANY_FRAMES) return fail(ctx, PCMB200_ERR_CAPACITY, "pcmb200_aggregator_push: more flushed frames than max_frames; push rejected, nothing changed");
    return fail(ctx, PCMB200_ERR_CAPACITY, "pcmb200_aggregator_push: flushed frames exceed the output capacity; push rejected, nothing changed");
}

template <typename T>
cudaError_t agg_alloc(T **p, size_t count)
{
    return cudaMalloc((void **)p, (count ? count : 1) * sizeof(T));
}

// One link of the push chain.  With tune.agg_pdl the launch carries the programmatic-serialisation attribute
// (see pdl_enter in agg_kernels.cuh); without it this is a plain <<<>>> launch.
template <typename... Params, typename... Args>
void agg_launch(pcmb200_ctx *ctx, void (*kern)(Params...), unsigned grid, unsigned block, size_t smem, Args &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = ctx->tune.agg_pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
    ctx->launches += 1;
}

// Enqueue the whole push on the context's stream.  Every pointer is device memory.
int agg_enqueue(pcmb200_ctx *ctx, pcmb200_aggregator *a, uint32_t n, const uint32_t *sessions, const double *arrival,
                const float *samples, int16_t *out, size_t out_capacity, uint32_t *frame_session, uint32_t *frame_len,
                unsigned long long *frame_offset, uint32_t max_frames)
{
    cudaStream_t st = ctx->stream;
    // Two control blocks, used in turn: the kernels of this push work on `ctl`, and the last one clears the other
    // block for the next push (that one was read back at the end of the previous push, which this stream has
    // finished).  Saves a memset call per push on a path that is bound by the host's enqueue time.
    AggCtl *ctl = a->ctl + a->ctl_idx, *ctl_next = a->ctl + (a->ctl_idx ^ 1);
    a->ctl_idx ^= 1;
    // kernels with one thread per QUANTUM, and kernels with one thread per TOUCHED SESSION (at most min(n, max_sessions))
    const uint32_t max_touched = std::min(n, a->geo.max_sessions);
    const unsigned blocks_n = (n + 255u) / 256u, blocks_t = (max_touched + 255u) / 256u, blocks_t128 = (max_touched + 127u) / 128u;
    const unsigned scan_blocks = (n + kScanBlock - 1) / kScanBlock;
    agg_launch(ctx, k_agg_count, blocks_n, 256, 0, sessions, n, a->geo.max_sessions, a->count, a->touched, a->touched_idx, ctl);
    ScanArgs sa{n, a->fl_nq, a->nout_of, ctl};
    agg_launch(ctx, k_agg_reserve, blocks_t, 256, 0, a->touched, a->count, a->cursor, ctl);
    agg_launch(ctx, k_agg_scatter, blocks_n, 256, 0, sessions, n, a->cursor, a->sorted, ctl);
    AggWalkOut wo{a->fl_nq, a->fl_ring, a->fl_pos, a->dst_slot, a->new_fill, a->flushed, a->new_last};
    // which build of the walk kernel: index lists of up to 32 quanta in registers (640 threads per SM), or up to 4 (2.5x
    // the threads) when the previous push averaged at most two quanta per touched session -- the streaming regime.
    // h_ctl still holds the previous push's record (it was waited for: one push may be pending).  Both are correct for
    // any push; a wrong guess costs time only.
    const bool short_lists = ctx->tune.agg_walk_regs == 4 ||
        (ctx->tune.agg_walk_regs == 0 && a->prev_n != 0 && a->h_ctl->n_touched != 0 && a->prev_n <= 2u * a->h_ctl->n_touched);
    a->prev_n = n;
    if (short_lists) agg_launch(ctx, k_agg_walk<4>, blocks_t128, 128, 0, a->geo, arrival, a->touched, a->touched_idx, a->count, a->cursor, a->sorted, a->fill, a->last_send, wo, ctl);
    else agg_launch(ctx, k_agg_walk<32>, blocks_t128, 128, 0, a->geo, arrival, a->touched, a->touched_idx, a->count, a->cursor, a->sorted, a->fill, a->last_send, wo, ctl);
    ScanApplyB ab{sessions, a->fl_ring, a->fl_pos, frame_session, frame_len, frame_offset, a->fdesc, a->fdesc_off, max_frames};
    agg_launch(ctx, k_scan_reduce, scan_blocks, kScanThreads, 0, sa, a->sums_a, a->sums_b);
    agg_launch(ctx, k_scan_apply, scan_blocks, kScanThreads, 0, sa, a->sums_a, a->sums_b, ab, ctl, (unsigned long long)out_capacity);

    AggConvertArgs ca{};
    ca.samples = samples; ca.staged = a->staged; ca.out_lo = a->out_lo; ca.nout_pad = a->nout_pad; ca.sorted = a->sorted; ca.fdesc = a->fdesc; ca.fdesc_off = a->fdesc_off;
    ca.nout_of = a->nout_of; ca.lut = a->lut; ca.out = out; ca.rms_at = a->rms_at; ca.ctl = ctl; ca.G = a->G;
    ca.magic_q = magic_for(a->geo.quantum); ca.out_aligned16 = aligned(out, 16) ? 1 : 0; ca.hint = ctx->tune.store_hint;
    // frames <= quanta; a frame has about desired_ms worth of quanta, so n / G blocks is an upper bound
    // that is far too generous: cap the grid and let the blocks loop
    unsigned long long conv_blocks = ((unsigned long long)n + a->G - 1) / a->G;
    const unsigned long long cap = (unsigned long long)ctx->sm_count * 32ull;
    if (conv_blocks > cap) conv_blocks = cap;
    const unsigned conv_threads = (unsigned)ctx->tune.agg_block;
    if (short_lists && ctx->tune.agg_warp_frames) {
        // the previous push had about one quantum per session: frames will be mostly carried-over outputs plus a
        // quantum or two -- a warp per frame (correct for any frame, fast for those)
        auto kern = a->meter ? (a->lut ? k_agg_convert_warp<true, true> : k_agg_convert_warp<true, false>)
                             : (a->lut ? k_agg_convert_warp<false, true> : k_agg_convert_warp<false, false>);
        unsigned &resident = a->warp_grid[a->meter ? 1 : 0];
        if (resident == 0) {
            int per_sm = 0;
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, a->convert_warp_smem));
            resident = (unsigned)(std::max(per_sm, 1) * ctx->sm_count);
        }
        const unsigned long long want = ((unsigned long long)n + 3) / 4;        // four frames per CTA trip, at most n frames
        agg_launch(ctx, kern, (unsigned)std::min<unsigned long long>(want, resident), 128u, a->convert_warp_smem, a->geo, ca);
    } else if (a->G == 1 && a->geo.half_quanta <= 64 && conv_threads <= 128) {
        // one frame per step and at most 64 quanta per frame: descriptors in registers, no meta buffers
        auto kern = a->meter ? (a->lut ? k_agg_convert_one<true, true> : k_agg_convert_one<true, false>)
                             : (a->lut ? k_agg_convert_one<false, true> : k_agg_convert_one<false, false>);
        // a persistent grid: exactly the CTAs that are resident at once, each walking over frames with the grid's
        // stride (a larger grid left the last wave two thirds full: 55 % of the warp slots active, ncu)
        const int key = (int)conv_threads * 2 + (a->meter ? 1 : 0);
        if (a->one_grid_key != key) {
            int per_sm = 0;
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, (int)conv_threads, a->convert_one_smem));
            a->one_grid = (unsigned)(std::max(per_sm, 1) * ctx->sm_count);
            a->one_grid_key = key;
        }
        if (ctx->tune.agg_persistent && conv_blocks > a->one_grid) conv_blocks = a->one_grid;
        agg_launch(ctx, kern, (unsigned)conv_blocks, conv_threads, a->convert_one_smem, a->geo, ca);
    } else {
        auto kern = a->meter ? (a->lut ? k_agg_convert_pipe<true, true> : k_agg_convert_pipe<true, false>)
                             : (a->lut ? k_agg_convert_pipe<false, true> : k_agg_convert_pipe<false, false>);
        agg_launch(ctx, kern, (unsigned)conv_blocks, conv_threads, a->convert_pipe_smem, a->geo, ca);
    }
    const AggAppendArgs apd{a->staged, a->out_lo, a->lut, a->nout_pad, a->geo.half_quanta, a->geo.stride, a->pairs, magic_for(a->pairs)};
    const AggCommitArgs cm{a->touched, a->count, a->fill, a->last_send, a->new_fill, a->new_last, ctl_next};
    if (a->meter)
        agg_launch(ctx, k_agg_levels, blocks_t128, 128, 0, a->geo, a->touched, a->count, a->cursor, a->fill, a->flushed, a->ring_rms, a->rms_at,
                   a->level, ctl);
    {
        // 64 quanta per 256-thread block step, four steps per trip when most quanta of the push are expected to flush
        // (trips with nothing to do then cost a quarter) and one when most are expected to wait (streaming: the previous
        // push had short lists; measured 3 % faster that way); a push small enough to fit one wave of the smaller blocks
        // takes 32 per 128, one step per trip (parallelism matters more than round trips there)
        const bool small = n <= 32u * 12u * (uint32_t)ctx->sm_count;
        const unsigned per_trip = small ? 32u : short_lists ? 64u : 256u;
        const unsigned long long trips = ((unsigned long long)n + per_trip - 1) / per_trip;
        const unsigned grid = (unsigned)(trips < cap ? trips : cap);
        const size_t smem = small ? a->append_smem / 2 : a->append_smem;
        if (a->meter) {
            if (small) agg_launch(ctx, k_agg_append<true, 32, 1>, grid, 128, smem, samples, a->dst_slot, apd, a->ring_rms, n, a->geo.quantum, ctl, cm);
            else if (short_lists) agg_launch(ctx, k_agg_append<true, 64, 1>, grid, 256, smem, samples, a->dst_slot, apd, a->ring_rms, n, a->geo.quantum, ctl, cm);
            else agg_launch(ctx, k_agg_append<true, 64, 4>, grid, 256, smem, samples, a->dst_slot, apd, a->ring_rms, n, a->geo.quantum, ctl, cm);
        } else {
            if (small) agg_launch(ctx, k_agg_append<false, 32, 1>, grid, 128, smem, samples, a->dst_slot, apd, a->ring_rms, n, a->geo.quantum, ctl, cm);
            else if (short_lists) agg_launch(ctx, k_agg_append<false, 64, 1>, grid, 256, smem, samples, a->dst_slot, apd, a->ring_rms, n, a->geo.quantum, ctl, cm);
            else agg_launch(ctx, k_agg_append<false, 64, 4>, grid, 256, smem, samples, a->dst_slot, apd, a->ring_rms, n, a->geo.quantum, ctl, cm);
        }
    }
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(a->h_ctl, ctl, sizeof(AggCtl), cudaMemcpyDeviceToHost, st));
    a->pending = true;
    return PCMB200_OK;
}

}  // namespace

extern "C" int pcmb200_aggregator_create(pcmb200_ctx *ctx, uint32_t max_sessions, double in_rate, uint32_t quantum,
                                         double desired_ms, uint32_t max_frame_quanta, uint32_t max_push_quanta,
                                         size_t max_out_samples_per_push, pcmb200_aggregator **out)
{
    NEED(ctx && out, "pcmb200_aggregator_create: NULL argument");
    *out = nullptr;
    NEED(max_sessions > 0 && quantum > 0 && max_frame_quanta > 0 && max_push_quanta > 0 && max_out_samples_per_push > 0,
         "pcmb200_aggregator_create: sizes must be positive");
    NEED(in_rate > 0 && desired_ms == desired_ms, "pcmb200_aggregator_create: bad rate / desired_ms");
    NEED(quantum % 4 == 0, "pcmb200_aggregator_create: quantum must be a multiple of 4 samples (the AudioWorklet quantum is 128)");
    NEED((unsigned long long)max_sessions * max_frame_quanta < 0xFFFFFFFFull, "pcmb200_aggregator_create: max_sessions * max_frame_quanta must be below 2^32");
    NEED((unsigned long long)max_frame_quanta * quantum * quantum < 0xF0000000ull, "pcmb200_aggregator_create: frame too long");
    size_t probe = 0;
    int s = resample_len((size_t)max_frame_quanta * quantum, in_rate, &probe);
    if (s) return fail(ctx, s, "pcmb200_aggregator_create: RangeError: invalid typed array length");
    NEED(probe <= 0x7FFFFFFFull, "pcmb200_aggregator_create: frame too long");
    // one frame (max_frame_quanta quanta at a pitch of quantum + 4 floats) must fit in shared memory
    const size_t slot_bytes = (size_t)(quantum + 4) * 4;
    NEED((size_t)max_frame_quanta * (slot_bytes + 8) <= 200 * 1024,
         "pcmb200_aggregator_create: max_frame_quanta * quantum is too long to stage one frame on chip (limit ~200 KB, "
         "e.g. 387 quanta of 128 samples = 1 s at 48 kHz)");
    // the append kernel brings 64 quanta at a time into shared memory (and the warp-per-frame convert kernel 16)
    NEED((size_t)64 * slot_bytes <= 220 * 1024, "pcmb200_aggregator_create: quantum too long (at most 876 samples; the AudioWorklet quantum is 128)");
    CU(cudaSetDevice(ctx->device));
    pcmb200_aggregator *a = new (std::nothrow) pcmb200_aggregator();
    if (!a) return fail(ctx, PCMB200_ERR_NOMEM, "pcmb200_aggregator_create: host allocation failed");
    a->geo.max_sessions = max_sessions; a->geo.quantum = quantum; a->geo.half_quanta = max_frame_quanta;
    a->geo.desired_ms = desired_ms;
    a->max_push = max_push_quanta; a->out_cap = max_out_samples_per_push; a->rate = in_rate;
    uint32_t R = 0;
    a->geo.stride = (in_rate == 16000.0) ? 1u : (integer_ratio(in_rate, &R) ? R : 0u);
    a->nout_max = (uint32_t)probe;

    // Frame arithmetic for every possible frame length, once.  The pick sequence floor(idx), idx += ratio
    // (App.tsx:26-30) does not depend on the input length -- a shorter frame uses a prefix of it -- so ONE
    // table serves every frame length; the per-length part is only the output length (App.tsx:24).
    std::vector<uint32_t> h_nout, h_lut, h_out_lo;
    try {
        h_nout.assign(max_frame_quanta + 1, 0);
        for (uint32_t nq = 1; nq <= max_frame_quanta; nq++) {
            size_t no = 0;
            resample_len((size_t)nq * quantum, in_rate, &no);
            h_nout[nq] = (uint32_t)no;
        }
        if (a->geo.stride == 0) {
            h_lut.assign((size_t)a->nout_max + 1, 0xFFFFFFFFu);        // one entry of padding: the convert kernel reads the table in pairs
            build_index_table((size_t)max_frame_quanta * quantum, in_rate, a->nout_max, h_lut.data());
        }
        // out_lo[j]: the first output whose pick index is >= j * quantum (picks never decrease with the output index),
        // i.e. quantum j of a frame owns the outputs [out_lo[j], out_lo[j + 1]) -- whatever the frame's final length
        h_out_lo.assign((size_t)max_frame_quanta + 1, 0);
        for (uint32_t j = 0, o = 0; j <= max_frame_quanta; j++) {
            const unsigned long long lim = (unsigned long long)j * quantum;
            while (o < a->nout_max && (a->geo.stride ? (unsigned long long)o * a->geo.stride : (unsigned long long)h_lut[o]) < lim) o++;
            h_out_lo[j] = o;
        }
        // the append kernel deals out aligned output pairs: the most any quantum position can touch
        a->pairs = 2;
        for (uint32_t j = 0; j < max_frame_quanta; j++)
            if (h_out_lo[j + 1] > h_out_lo[j])
                a->pairs = std::max(a->pairs, ((h_out_lo[j + 1] + 1u) / 2u) - (h_out_lo[j] / 2u));
    } catch (const std::bad_alloc &) { delete a; return fail(ctx, PCMB200_ERR_NOMEM, "pcmb200_aggregator_create: host allocation failed"); }

    // frames per convert step: as many whole worst-case frames as fit in ~64 KB, at least one
    a->G = (uint32_t)std::max<size_t>(1, std::min<size_t>(8, ((size_t)ctx->tune.agg_stage_kb * 1024) / ((size_t)max_frame_quanta * slot_bytes)));
    a->convert_pipe_smem = agg_arena_bytes(a->G, max_frame_quanta, quantum) + 2 * (size_t)agg_meta_layout(a->G, a->G * max_frame_quanta).bytes;
    a->convert_one_smem = agg_arena_bytes(1, max_frame_quanta, quantum);
    a->convert_warp_smem = (size_t)4 * kWarpWindow * slot_bytes;               // four warps, one window each
    a->append_smem = (((size_t)64 * slot_bytes + 127) & ~(size_t)127);

    cudaError_t e = cudaSuccess;
    const size_t S = max_sessions, P = max_push_quanta;
    a->nout_pad = (a->nout_max + 7u) & ~7u;
#define AGG_STEP(call) if (e == cudaSuccess) e = (call)
    AGG_STEP(agg_alloc(&a->staged, S * (size_t)a->nout_pad));
    AGG_STEP(agg_alloc(&a->out_lo, (size_t)max_frame_quanta + 1));
    AGG_STEP(agg_alloc(&a->ring_rms, S * max_frame_quanta));
    AGG_STEP(agg_alloc(&a->fill, S));
    AGG_STEP(agg_alloc(&a->count, S));
    AGG_STEP(agg_alloc(&a->cursor, S));
    AGG_STEP(agg_alloc(&a->last_send, S));
    AGG_STEP(agg_alloc(&a->level, S));
    AGG_STEP(agg_alloc(&a->nout_of, (size_t)max_frame_quanta + 1));
    if (a->geo.stride == 0) AGG_STEP(agg_alloc(&a->lut, h_lut.size()));
    AGG_STEP(agg_alloc(&a->d_sessions, P));
    AGG_STEP(agg_alloc(&a->d_arrival, P));
    // d_stage (a whole push of samples) is allocated on the first push whose samples are host memory
    AGG_STEP(agg_alloc(&a->touched, P));
    AGG_STEP(agg_alloc(&a->touched_idx, P));
    AGG_STEP(agg_alloc(&a->sorted, P));
    AGG_STEP(agg_alloc(&a->fl_nq, P));
    AGG_STEP(agg_alloc(&a->fl_ring, P));
    AGG_STEP(agg_alloc(&a->fl_pos, P));
    AGG_STEP(agg_alloc(&a->dst_slot, P));
    AGG_STEP(agg_alloc(&a->new_fill, P));
    AGG_STEP(agg_alloc(&a->flushed, P));
    AGG_STEP(agg_alloc(&a->new_last, P));
    AGG_STEP(agg_alloc(&a->rms_at, P));
    const size_t nb = (P + kScanBlock - 1) / kScanBlock;
    AGG_STEP(agg_alloc(&a->sums_a, nb));
    AGG_STEP(agg_alloc(&a->sums_b, nb));
    AGG_STEP(agg_alloc(&a->fdesc, P));
    AGG_STEP(agg_alloc(&a->fdesc_off, P));
    AGG_STEP(agg_alloc(&a->f_session, P));
    AGG_STEP(agg_alloc(&a->f_len, P));
    AGG_STEP(agg_alloc(&a->f_off, P));
    AGG_STEP(agg_alloc(&a->d_out, max_out_samples_per_push));
    AGG_STEP(agg_alloc(&a->ctl, 2));
    AGG_STEP(cudaMemsetAsync(a->ctl, 0, 2 * sizeof(AggCtl), ctx->stream));
    AGG_STEP(cudaHostAlloc((void **)&a->h_ctl, sizeof(AggCtl), cudaHostAllocDefault));
    AGG_STEP(cudaMemsetAsync(a->fill, 0, S * 4, ctx->stream));
    AGG_STEP(cudaMemsetAsync(a->count, 0, S * 4, ctx->stream));
    AGG_STEP(cudaMemsetAsync(a->last_send, 0, S * 8, ctx->stream));
    AGG_STEP(cudaMemsetAsync(a->level, 0, S * 8, ctx->stream));
    AGG_STEP(cudaMemsetAsync(a->staged, 0, S * (size_t)a->nout_pad * 2, ctx->stream));
    AGG_STEP(cudaMemcpyAsync(a->out_lo, h_out_lo.data(), h_out_lo.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    AGG_STEP(cudaMemsetAsync(a->ring_rms, 0, S * max_frame_quanta * 8, ctx->stream));
    AGG_STEP(cudaMemcpyAsync(a->nout_of, h_nout.data(), h_nout.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (a->geo.stride == 0) AGG_STEP(cudaMemcpyAsync(a->lut, h_lut.data(), h_lut.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_pipe<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_pipe_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_pipe<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_pipe_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_pipe<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_pipe_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_pipe<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_pipe_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_warp<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_warp_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_warp<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_warp_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_warp<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_warp_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_convert_warp<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_warp_smem));
    if (a->G == 1 && max_frame_quanta <= 64) {
        AGG_STEP(cudaFuncSetAttribute(k_agg_convert_one<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_one_smem));
        AGG_STEP(cudaFuncSetAttribute(k_agg_convert_one<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_one_smem));
        AGG_STEP(cudaFuncSetAttribute(k_agg_convert_one<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_one_smem));
        AGG_STEP(cudaFuncSetAttribute(k_agg_convert_one<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->convert_one_smem));
    }
    AGG_STEP(cudaFuncSetAttribute(k_agg_append<true, 64, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->append_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_append<false, 64, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->append_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_append<true, 64, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->append_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_append<false, 64, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->append_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_append<true, 32, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->append_smem));
    AGG_STEP(cudaFuncSetAttribute(k_agg_append<false, 32, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a->append_smem));
    AGG_STEP(cudaStreamSynchronize(ctx->stream));     // the host vectors above go out of scope
#undef AGG_STEP
    if (e != cudaSuccess) {
        std::string m = std::string("pcmb200_aggregator_create: ") + cudaGetErrorString(e);
        pcmb200_aggregator_destroy(ctx, a);
        cudaGetLastError();
        return fail(ctx, e == cudaErrorMemoryAllocation ? PCMB200_ERR_NOMEM : PCMB200_ERR_CUDA, m);
    }
    if (a->h_ctl) std::memset(a->h_ctl, 0, sizeof(AggCtl));
    *out = a;
    return PCMB200_OK;
}

extern "C" int pcmb200_aggregator_destroy(pcmb200_ctx *ctx, pcmb200_aggregator *a)
{
    NEED(ctx, "pcmb200_aggregator_destroy: ctx is NULL");
    if (!a) return PCMB200_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    void *dev[] = {a->touched_idx, a->staged, a->out_lo, a->ring_rms, a->fill, a->count, a->cursor, a->last_send, a->level, a->nout_of, a->lut, a->d_sessions,
                   a->d_arrival, a->d_stage, a->touched, a->sorted, a->fl_nq, a->fl_ring, a->fl_pos, a->dst_slot, a->new_fill,
                   a->flushed, a->new_last, a->rms_at, a->sums_a, a->sums_b, a->fdesc, a->fdesc_off, a->f_session,
                   a->f_len, a->f_off, a->d_out, a->ctl};
    for (void *p : dev) if (p) cudaFree(p);
    if (a->h_ctl) cudaFreeHost(a->h_ctl);
    cudaGetLastError();
    delete a;
    return PCMB200_OK;
}

// App.tsx:325-330 (stop) / :148-157 (initial state): empty buffer, lastSend = now.  Enqueued on the
// context's stream, so it is ordered with the pushes.
extern "C" int pcmb200_aggregator_reset_session(pcmb200_ctx *ctx, pcmb200_aggregator *a, uint32_t session, double now_ms)
{
    NEED(ctx && a, "pcmb200_aggregator_reset_session: NULL argument");
    NEED(session < a->geo.max_sessions, "pcmb200_aggregator_reset_session: session out of range");
    CU(cudaSetDevice(ctx->device));
    k_agg_reset<<<1, 32, 0, ctx->stream>>>(a->fill, a->last_send, session, 1u, now_ms);
    CU(cudaGetLastError());
    return PCMB200_OK;
}

extern "C" int pcmb200_aggregator_reset_range(pcmb200_ctx *ctx, pcmb200_aggregator *a, uint32_t first_session, uint32_t count,
                                              double now_ms)
{
    NEED(ctx && a, "pcmb200_aggregator_reset_range: NULL argument");
    NEED((unsigned long long)first_session + count <= a->geo.max_sessions, "pcmb200_aggregator_reset_range: session range out of bounds");
    if (count == 0) return PCMB200_OK;
    CU(cudaSetDevice(ctx->device));
    k_agg_reset<<<(int)((count + 255u) / 256u), 256, 0, ctx->stream>>>(a->fill, a->last_send, first_session, count, now_ms);
    CU(cudaGetLastError());
    return PCMB200_OK;
}

// Level meter of the aggregated sessions (App.tsx:267-271).  When enabled, every quantum's rms is taken
// as the quantum passes through the chip, and the level runs through the quanta of every flushed frame
// in arrival order; the level of a session therefore lags the reference's by the quanta still waiting
// in its buffer (it is a UI-only value there).  Switch it on before the first push: quanta appended
// while it was off have no rms on record.
extern "C" int pcmb200_aggregator_set_meter(pcmb200_ctx *ctx, pcmb200_aggregator *a, int enabled)
{
    NEED(ctx && a, "pcmb200_aggregator_set_meter: NULL argument");
    a->meter = enabled != 0;
    return PCMB200_OK;
}

extern "C" int pcmb200_aggregator_get_levels(pcmb200_ctx *ctx, pcmb200_aggregator *a, uint32_t first_session, uint32_t count,
                                             double *levels_host)
{
    NEED(ctx && a && (levels_host || count == 0), "pcmb200_aggregator_get_levels: NULL argument");
    NEED((unsigned long long)first_session + count <= a->geo.max_sessions, "pcmb200_aggregator_get_levels: session range out of bounds");
    CU(cudaSetDevice(ctx->device));
    if (count) CU(cudaMemcpyAsync(levels_host, a->level + first_session, (size_t)count * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return PCMB200_OK;
}

// Everything on the device: quanta, their session ids and arrival times, the frames and the frame
// arrays.  Asynchronous; pcmb200_aggregator_wait collects the outcome.
extern "C" int pcmb200_aggregator_push_device(pcmb200_ctx *ctx, pcmb200_aggregator *a, uint32_t n_quanta,
                                              const uint32_t *sessions_device, const double *arrival_ms_device,
                                              const float *samples_device, int16_t *out_device, size_t out_capacity,
                                              uint32_t *frame_session_device, uint32_t *frame_len_device,
                                              unsigned long long *frame_offset_device, uint32_t max_frames)
{
    NEED(ctx && a, "pcmb200_aggregator_push_device: NULL argument");
    NEED(!a->pending, "pcmb200_aggregator_push_device: collect the previous push with pcmb200_aggregator_wait first");
    if (n_quanta == 0) { std::memset(a->h_ctl, 0, sizeof(AggCtl)); a->pending = true; return PCMB200_OK; }
    NEED(sessions_device && arrival_ms_device && samples_device && out_device, "pcmb200_aggregator_push_device: NULL buffer");
    NEED(n_quanta <= a->max_push, "pcmb200_aggregator_push_device: more quanta than max_push_quanta");
    NEED(aligned(samples_device, 16), "pcmb200_aggregator_push_device: samples must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    return agg_enqueue(ctx, a, n_quanta, sessions_device, arrival_ms_device, samples_device, out_device, out_capacity,
                       frame_session_device, frame_len_device, frame_offset_device, max_frames);
}

extern "C" int pcmb200_aggregator_wait(pcmb200_ctx *ctx, pcmb200_aggregator *a, uint32_t *n_frames, size_t *out_samples)
{
    NEED(ctx && a, "pcmb200_aggregator_wait: NULL argument");
    if (n_frames) *n_frames = 0;
    if (out_samples) *out_samples = 0;
    NEED(a->pending, "pcmb200_aggregator_wait: no push is pending");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    a->pending = false;
    if (a->h_ctl->status) return agg_status_to_error(ctx, a->h_ctl->status);
    if (n_frames) *n_frames = a->h_ctl->n_frames;
    if (out_samples) *out_samples = (size_t)a->h_ctl->out_total;
    return PCMB200_OK;
}

extern "C" int pcmb200_aggregator_push(pcmb200_ctx *ctx, pcmb200_aggregator *a, uint32_t n_quanta,
                                       const uint32_t *sessions, const double *arrival_ms, const float *samples,
                                       int samples_mem, int16_t *out, size_t out_capacity, int out_mem,
                                       uint32_t *frame_session, uint32_t *frame_len, unsigned long long *frame_offset,
                                       uint32_t max_frames, uint32_t *n_frames)
{
    NEED(ctx && a && n_frames, "pcmb200_aggregator_push: NULL argument");
    *n_frames = 0;
    NEED(!a->pending, "pcmb200_aggregator_push: collect the previous asynchronous push with pcmb200_aggregator_wait first");
    if (n_quanta == 0) return PCMB200_OK;
    NEED(sessions && arrival_ms && samples && out, "pcmb200_aggregator_push: NULL buffer");
    NEED((samples_mem == PCMB200_MEM_HOST || samples_mem == PCMB200_MEM_DEVICE) &&
         (out_mem == PCMB200_MEM_HOST || out_mem == PCMB200_MEM_DEVICE), "pcmb200_aggregator_push: bad mem kind");
    NEED(n_quanta <= a->max_push, "pcmb200_aggregator_push: more quanta than max_push_quanta");
    NEED(samples_mem == PCMB200_MEM_HOST || aligned(samples, 16), "pcmb200_aggregator_push: device samples must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    // the session ids and arrival times are host arrays in this entry point: stage them (pinned caller
    // memory goes at PCIe speed, pageable memory through the driver's bounce buffer)
    CU(cudaMemcpyAsync(a->d_sessions, sessions, (size_t)n_quanta * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(a->d_arrival, arrival_ms, (size_t)n_quanta * 8, cudaMemcpyHostToDevice, st));
    const float *d_samples = samples;
    if (samples_mem == PCMB200_MEM_HOST) {
        if (!a->d_stage) CU(agg_alloc(&a->d_stage, (size_t)a->max_push * a->geo.quantum));
        CU(cudaMemcpyAsync(a->d_stage, samples, (size_t)n_quanta * a->geo.quantum * sizeof(float), cudaMemcpyHostToDevice, st));
        d_samples = a->d_stage;
    }
    int16_t *d_out = out_mem == PCMB200_MEM_DEVICE ? out : a->d_out;
    const size_t cap = out_mem == PCMB200_MEM_DEVICE ? out_capacity : std::min(out_capacity, a->out_cap);
    if (max_frames > a->max_push) max_frames = a->max_push;       // frames <= quanta; the internal arrays hold max_push
    int s = agg_enqueue(ctx, a, n_quanta, a->d_sessions, a->d_arrival, d_samples, d_out, cap, a->f_session, a->f_len, a->f_off, max_frames);
    if (s) return s;
    CU(cudaStreamSynchronize(st));
    a->pending = false;
    if (a->h_ctl->status) return agg_status_to_error(ctx, a->h_ctl->status);
    const uint32_t nf = a->h_ctl->n_frames;
    if (nf) {
        if (frame_session) CU(cudaMemcpyAsync(frame_session, a->f_session, (size_t)nf * 4, cudaMemcpyDeviceToHost, st));
        if (frame_len) CU(cudaMemcpyAsync(frame_len, a->f_len, (size_t)nf * 4, cudaMemcpyDeviceToHost, st));
        if (frame_offset) CU(cudaMemcpyAsync(frame_offset, a->f_off, (size_t)nf * 8, cudaMemcpyDeviceToHost, st));
        // the frames themselves: out_total counts every frame padded to 8 samples, and the capacity check
        // was made on that padded total, so this copy stays inside both buffers
        if (out_mem == PCMB200_MEM_HOST && a->h_ctl->out_total)
            CU(cudaMemcpyAsync(out, a->d_out, (size_t)a->h_ctl->out_total * sizeof(int16_t), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    *n_frames = nf;
    return PCMB200_OK;
}
