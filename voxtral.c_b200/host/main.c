/*
 * voxtral_b200 - command line front end of the CUDA backend (the hot-path subset of the
 * reference CLI, /root/reference/main.c:121-411):
 *
 *   voxtral_b200 -d <model_dir> -i <file.wav> [-I <seconds>] [--stream <chunk_seconds>]
 *   voxtral_b200 -d <model_dir> --stdin [-I <seconds>]          (s16le, 16 kHz, mono on stdin)
 *
 * Prints the greedy token ids, one per line, on stdout (the tokenizer is outside the hot path)
 * and a timing summary on stderr in the spirit of the reference's (voxtral.c:1359-1370).
 * Without --stream/--stdin the clip goes through the one-call batch path (vox_transcribe_audio);
 * with them it is fed through the streaming API in chunks.
 */
#include "vox_host.h"
#include "vox_cuda.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

static double now_ms(void) {
    struct timeval t;
    gettimeofday(&t, NULL);
    return t.tv_sec * 1e3 + t.tv_usec * 1e-3;
}

static void usage(const char *p) {
    fprintf(stderr, "usage: %s -d <model_dir> (-i <file.wav> | --stdin) [-I <seconds>] [--stream <chunk_seconds>] [-v]\n", p);
}

int main(int argc, char **argv) {
    const char *model_dir = NULL, *input = NULL;
    float interval = 2.0f, feed_chunk = 0.0f;
    int from_stdin = 0, verbose = 0;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "-d") && i + 1 < argc) model_dir = argv[++i];
        else if (!strcmp(argv[i], "-i") && i + 1 < argc) input = argv[++i];
        else if (!strcmp(argv[i], "-I") && i + 1 < argc) interval = (float)atof(argv[++i]);
        else if (!strcmp(argv[i], "--stream") && i + 1 < argc) feed_chunk = (float)atof(argv[++i]);
        else if (!strcmp(argv[i], "--stdin")) from_stdin = 1;
        else if (!strcmp(argv[i], "-v")) verbose = 1;
        else { usage(argv[0]); return 2; }
    }
    if (!model_dir || (!input && !from_stdin)) { usage(argv[0]); return 2; }

    double t0 = now_ms();
    vox_b200_ctx *ctx = vox_b200_load(model_dir);
    if (!ctx) return 1;
    fprintf(stderr, "Model loaded in %.0f ms (%.2f GB on the device)\n", now_ms() - t0, vox_cuda_memory_used() / 1e9);

    float *samples = NULL;
    int n = 0;
    if (!from_stdin) {
        samples = vox_b200_load_wav(input, &n);
        if (!samples) { vox_b200_free(ctx); return 1; }
    }
    int rc = 0;
    if (!from_stdin && feed_chunk <= 0.0f) {
        const int max_ids = n / 1280 + 128;
        int *ids = (int *)malloc(sizeof(int) * (size_t)max_ids);
        float dev_ms = 0.f;
        t0 = now_ms();
        const int k = vox_b200_transcribe_ids(ctx, samples, n, ids, max_ids, &dev_ms);
        const double wall = now_ms() - t0;
        if (k < 0) rc = 1;
        for (int i = 0; i < k; i++) printf("%d\n", ids[i]);
        fprintf(stderr, "Batch: %.2f s of audio in %.1f ms (device %.1f ms), %d tokens, %.1fx real time\n",
                n / 16000.0, wall, dev_ms, k, (n / 16.0) / wall);
        free(ids);
    } else {
        vox_b200_stream *s = vox_b200_stream_init(ctx);
        if (!s) { vox_b200_free(ctx); return 1; }
        vox_b200_set_processing_interval(s, interval);
        int ids[64];
        t0 = now_ms();
        long fed = 0;
        if (from_stdin) {
            int16_t raw[4096];
            float f[4096];
            size_t got;
            while ((got = fread(raw, sizeof(int16_t), 4096, stdin)) > 0) {
                for (size_t i = 0; i < got; i++) f[i] = (float)raw[i] / 32768.0f;
                if (vox_b200_stream_feed(s, f, (int)got)) { rc = 1; break; }
                fed += (long)got;
                int k;
                while ((k = vox_b200_stream_get_ids(s, ids, 64)) > 0)
                    for (int i = 0; i < k; i++) printf("%d\n", ids[i]);
                fflush(stdout);
            }
        } else {
            const int chunk = (int)(feed_chunk * 16000.0f) > 0 ? (int)(feed_chunk * 16000.0f) : n;
            for (int off = 0; off < n && !rc; off += chunk) {
                const int c = n - off < chunk ? n - off : chunk;
                if (vox_b200_stream_feed(s, samples + off, c)) rc = 1;
                fed += c;
                int k;
                while ((k = vox_b200_stream_get_ids(s, ids, 64)) > 0)
                    for (int i = 0; i < k; i++) printf("%d\n", ids[i]);
            }
        }
        if (!rc && vox_b200_stream_finish(s)) rc = 1;
        int k;
        while ((k = vox_b200_stream_get_ids(s, ids, 64)) > 0)
            for (int i = 0; i < k; i++) printf("%d\n", ids[i]);
        const double wall = now_ms() - t0;
        double mel_ms, enc_ms, dec_ms;
        int calls, steps;
        vox_b200_stream_stats(s, &mel_ms, &enc_ms, &dec_ms, &calls, &steps);
        fprintf(stderr, "Audio: %ld samples (%.1f seconds)\n", fed, fed / 16000.0);     /* voxtral_audio.c summary line */
        vox_b200_stream_report(s, stderr, verbose);
        fprintf(stderr, "Mel: %.1f ms; encoder %.2f ms/call over %d calls; %d decoder steps\n"
                        "Stream: %.2f s of audio in %.1f ms, %.1fx real time\n",
                mel_ms, calls ? enc_ms / calls : 0.0, calls, steps, fed / 16000.0, wall, (fed / 16.0) / wall);
        vox_b200_stream_free(s);
    }
    free(samples);
    vox_b200_free(ctx);
    vox_cuda_shutdown();
    return rc;
}
