/* frames_c_abi.c -- the C ABI from plain C: convert one batch of 48 kHz frames held in host memory.
 *
 *   gcc -std=c11 -I include voice-enabled-browser-automation_b200/examples/frames_c_abi.c \
 *       -L voice-enabled-browser-automation_b200/lib -lpcm_b200 -Wl,-rpath,'$ORIGIN/../lib' -lm -o frames_c_abi
 *
 * This is the call sequence an FFI binding (N-API, cgo, JNI, ctypes) makes; see INTEGRATION.md.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "pcm_b200.h"

int main(void)
{
    printf("libpcm_b200 ABI %d, wire little-endian: %d\n", pcmb200_version(), pcmb200_wire_is_little_endian());

    /* frame arithmetic needs no device: floor(960 / 3) = 320, and 23 quanta -> 981 (App.tsx:23-24) */
    size_t n_out = 0;
    if (pcmb200_resample_len(2944, 48000.0, &n_out) != PCMB200_OK || n_out != 981) return 2;
    if (pcmb200_resample_len(960, 48000.0, &n_out) != PCMB200_OK || n_out != 320) return 2;
    if (pcmb200_resample_len(960, 0.0, &n_out) != PCMB200_ERR_RANGE) return 2;   /* JS throws RangeError */

    pcmb200_ctx *ctx = NULL;
    int st = pcmb200_ctx_create(0, &ctx);
    if (st != PCMB200_OK) {   /* no CPU fallback: report and stop */
        printf("no device: %s\n", pcmb200_last_error(NULL));
        return st == PCMB200_ERR_NO_DEVICE ? 0 : 1;
    }
    enum { ROWS = 64, N_IN = 960 };
    float *in = (float *)malloc(sizeof(float) * ROWS * N_IN);
    int16_t *out = (int16_t *)malloc(sizeof(int16_t) * ROWS * 320);
    for (int r = 0; r < ROWS; r++)
        for (int k = 0; k < N_IN; k++) in[r * N_IN + k] = 0.8f * sinf(6.2831853f * (110.0f + r) * k / 48000.0f);
    st = pcmb200_frames_encode(ctx, in, N_IN, out, 320, ROWS, N_IN, 48000.0, PCMB200_MEM_HOST, &n_out);
    if (st != PCMB200_OK) { printf("error: %s\n", pcmb200_last_error(ctx)); return 1; }
    printf("%d frames of %zu samples, first samples: %d %d %d\n", ROWS, n_out, out[0], out[1], out[2]);
    free(in);
    free(out);
    return pcmb200_ctx_destroy(ctx);
}
