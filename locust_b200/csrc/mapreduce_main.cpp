// MapReduce -- drop-in for the reference binary's single-mode entry point
//   ./MapReduce input_file.txt
//   mapreduce <filename> [line_start] [line_end] [node_num] [stage]
// (MapReduce/src/main.cu:358-532, usage string :364).  Host side only: argv, file
// read, stdout.  All computing happens in liblocust.so on the GPU.
//
// stdout keeps the reference's framing byte for byte: "Running", "Length: %i",
// the three stage-time lines (:408, :418, :468; here device-event times), one
// "print key:" line per word (:132), "\nDone".  LOCUST_FORMAT=cpu switches the key
// lines to the CPU branch's format (:286), which is what the parity oracle prints.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>
#include "locust.h"

static int die(locust_ctx *ctx, const char *what, int rc) {
    fprintf(stderr, "MapReduce: %s failed (%d): %s\n", what, rc, ctx ? locust_last_error(ctx) : "");
    if (ctx) locust_destroy(ctx);
    return 1;
}

int main(int argc, char *argv[]) {
    printf("Running\n");
    if (argc < 2) {
        printf("Missing or invalid arguments.\n");
        printf("mapreduce <filename> [line_start] [line_end] [node_num] [stage]\n");
        return -1;
    }
    long start_line = -1, end_line = -1;
    if (argc > 2) {
        if (argc < 4) {   // the reference reads argv[3] unchecked (main.cu:372) and crashes
            printf("Missing or invalid arguments.\n");
            printf("mapreduce <filename> [line_start] [line_end] [node_num] [stage]\n");
            return -1;
        }
        start_line = strtol(argv[2], nullptr, 10);
        end_line = strtol(argv[3], nullptr, 10);
        printf("Using custom start and end locations: (%i, %i)\n", (int)start_line, (int)end_line);
    }
    int stage = 0;
    if (argc > 4) {
        if (argc < 6) {   // main.cu:381 reads argv[5] unchecked
            printf("Missing or invalid arguments.\n");
            printf("mapreduce <filename> [line_start] [line_end] [node_num] [stage]\n");
            return -1;
        }
        stage = (int)strtol(argv[5], nullptr, 10);
    }
    if (stage != 0) {
        fprintf(stderr, "MapReduce: multi mode (stage %d) hands partials over with the NCCL shuffle of liblocust, "
                        "not through /tmp/out.txt; run single mode.\n", stage);
        return 2;
    }

    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 1; }
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t *buf = nullptr;
    if (cudaMallocHost(reinterpret_cast<void **>(&buf), sz > 0 ? (size_t)sz : 1) != cudaSuccess) {
        fprintf(stderr, "MapReduce: no CUDA device / pinned allocation failed (this build has no CPU path)\n");
        return 1;
    }
    if (sz > 0 && fread(buf, 1, (size_t)sz, f) != (size_t)sz) { fprintf(stderr, "MapReduce: short read\n"); return 1; }
    fclose(f);

    // loadFile (main.cu:40-64): the [line_start, line_end) window, by line index
    size_t lo = 0, hi = (size_t)sz;
    long nlines = 0;
    {
        long line = 0;
        size_t i = 0;
        bool lo_set = start_line < 0, hi_set = false;
        if (lo_set) lo = 0;
        while (i < (size_t)sz) {
            if (!lo_set && line >= start_line) { lo = i; lo_set = true; }
            if (start_line >= 0 && line >= end_line) { hi = i; hi_set = true; break; }
            const void *nl = memchr(buf + i, '\n', (size_t)sz - i);
            const size_t e = nl ? (size_t)((const uint8_t *)nl - buf) + 1 : (size_t)sz;
            if (lo_set) ++nlines;
            i = e;
            ++line;
        }
        if (!lo_set) { lo = (size_t)sz; }
        if (!hi_set) hi = (size_t)sz;
        if (hi < lo) hi = lo;
    }
    if (memchr(buf + lo, 0, hi - lo)) {
        fprintf(stderr, "MapReduce: input contains NUL bytes; the reference stops a line at NUL, this build refuses\n");
        return 1;
    }
    printf("Length: %i\n", (int)nlines);

    locust_config cfg;
    memset(&cfg, 0, sizeof cfg);
    if (const char *d = getenv("LOCUST_DEVICE")) cfg.device = atoi(d);
    if (const char *t = getenv("LOCUST_TABLE_LOG2")) cfg.table_log2 = (uint32_t)atoi(t);
    locust_ctx *ctx = nullptr;
    int rc = locust_create(&cfg, &ctx);
    if (rc) return die(ctx, "locust_create", rc);

    rc = locust_map(ctx, buf + lo, hi - lo, LOCUST_HOST);
    if (rc) return die(ctx, "locust_map", rc);
    rc = locust_process(ctx);
    if (rc) return die(ctx, "locust_process", rc);
    rc = locust_reduce(ctx);
    if (rc) return die(ctx, "locust_reduce", rc);
    printf("GPU mapping %d nanoseconds \n", (int)(locust_stage_ms(ctx, LOCUST_STAGE_MAP) * 1e6f));
    printf("GPU stream compaction and sorting %d nanoseconds \n", (int)(locust_stage_ms(ctx, LOCUST_STAGE_PROCESS) * 1e6f));
    printf("GPU reduce %d nanoseconds \n", (int)(locust_stage_ms(ctx, LOCUST_STAGE_REDUCE) * 1e6f));

    const char *fmt = getenv("LOCUST_FORMAT");
    const char *text = nullptr;
    uint64_t nbytes = 0;
    rc = locust_format_result(ctx, fmt && strcmp(fmt, "cpu") == 0, &text, &nbytes);
    if (rc) return die(ctx, "locust_format_result", rc);
    fwrite(text, 1, nbytes, stdout);
    locust_destroy(ctx);
    cudaFreeHost(buf);
    printf("\nDone\n");
    return 0;
}
