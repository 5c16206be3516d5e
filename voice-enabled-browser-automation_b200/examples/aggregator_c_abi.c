/* aggregator_c_abi.c -- the capture aggregator (App.tsx:148-157, :273-288) from plain C: three connections forward
 * 128-sample quanta, the library decides when each buffer flushes and returns the frames.
 *
 *   gcc -std=c11 -I include voice-enabled-browser-automation_b200/examples/aggregator_c_abi.c \
 *       -L voice-enabled-browser-automation_b200/lib -lpcm_b200 -Wl,-rpath,'$ORIGIN/../lib' -lm -o aggregator_c_abi
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "pcm_b200.h"

int main(void)
{
    pcmb200_ctx *ctx = NULL;
    int st = pcmb200_ctx_create(0, &ctx);
    if (st != PCMB200_OK) {   /* no CPU fallback: report and stop */
        printf("no device: %s\n", pcmb200_last_error(NULL));
        return st == PCMB200_ERR_NO_DEVICE ? 0 : 1;
    }
    enum { SESSIONS = 3, QUANTUM = 128, TICKS = 30 };
    pcmb200_aggregator *agg = NULL;
    /* 48 kHz capture, the reference's 60 ms, frames of at most 32 quanta, pushes of at most 64 quanta */
    st = pcmb200_aggregator_create(ctx, SESSIONS, 48000.0, QUANTUM, 60.0, 32, 64, 8192, &agg);
    if (st != PCMB200_OK) { printf("error: %s\n", pcmb200_last_error(ctx)); return 1; }
    pcmb200_aggregator_reset_range(ctx, agg, 0, SESSIONS, 0.0);       /* lastSend = 0 for every connection */

    uint32_t sessions[SESSIONS], frame_session[SESSIONS], frame_len[SESSIONS], n_frames = 0;
    unsigned long long frame_offset[SESSIONS];
    double arrival[SESSIONS];
    float *samples = (float *)malloc(sizeof(float) * SESSIONS * QUANTUM);
    int16_t *out = (int16_t *)malloc(sizeof(int16_t) * 8192);
    int total = 0;
    for (int q = 0; q < TICKS; q++) {                                   /* one quantum per connection per push */
        for (int s = 0; s < SESSIONS; s++) {
            sessions[s] = (uint32_t)s;
            arrival[s] = (q + 1) * (QUANTUM * 1e3 / 48000.0) + 0.1 * s;   /* this quantum's performance.now() */
            for (int k = 0; k < QUANTUM; k++)
                samples[s * QUANTUM + k] = 0.5f * sinf(6.2831853f * (220.0f + 110.0f * s) * (q * QUANTUM + k) / 48000.0f);
        }
        st = pcmb200_aggregator_push(ctx, agg, SESSIONS, sessions, arrival, samples, PCMB200_MEM_HOST, out, 8192, PCMB200_MEM_HOST,
                                     frame_session, frame_len, frame_offset, SESSIONS, &n_frames);
        if (st != PCMB200_OK) { printf("error: %s\n", pcmb200_last_error(ctx)); return 1; }
        for (uint32_t f = 0; f < n_frames; f++) {
            printf("quantum %2d: connection %u flushed %u samples (first: %d)\n", q, frame_session[f], frame_len[f],
                   out[frame_offset[f]]);
            if (frame_len[f] != 981) return 3;                          /* 23 quanta = 2944 samples -> 981 at 16 kHz */
            total++;
        }
    }
    printf("%d frames from %d connections\n", total, SESSIONS);
    free(samples);
    free(out);
    pcmb200_aggregator_destroy(ctx, agg);
    return total == SESSIONS ? pcmb200_ctx_destroy(ctx) : 4;
}
