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
//
// Multi mode (main.cu:421-446): `stage` 1 maps the line window and writes the combined
// (word, count) records as "%s \t%d\n" lines to /tmp/out.txt (LOCUST_INTERMEDIATE overrides
// the path); `stage` 2 reads such a file -- also one written by the reference's own stage 1,
// which holds one line per token -- and reduces it.  Inside one box the NCCL shuffle of
// liblocust replaces this file hand-off; the stages are kept so the binary still fits the
// reference's launcher (Distributor/slave.py starts it with these arguments).
//
// LOCUST_DEVICES=0,1,2,...  (single mode) runs the job on several GPUs of the box through
// locust_multi_*: the file is cut on delimiters, one shard per listed device, and the records are
// exchanged over NVLink -- what the reference's launcher + line windows + /tmp/out.txt do across
// machines.  The stdout is the same as on one GPU.
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
    if (stage != 0 && stage != 1 && stage != 2) {
        fprintf(stderr, "MapReduce: stage must be 0 (single), 1 (map) or 2 (reduce)\n");
        return 2;
    }
    const char *inter_path = getenv("LOCUST_INTERMEDIATE") ? getenv("LOCUST_INTERMEDIATE") : "/tmp/out.txt";
    const char *fmt_env = getenv("LOCUST_FORMAT");
    const int key_format = (fmt_env && strcmp(fmt_env, "cpu") == 0) ? LOCUST_FORMAT_CPU : LOCUST_FORMAT_GPU;

    locust_config cfg;
    memset(&cfg, 0, sizeof cfg);
    if (const char *d = getenv("LOCUST_DEVICE")) cfg.device = atoi(d);
    if (const char *t = getenv("LOCUST_TABLE_LOG2")) cfg.table_log2 = (uint32_t)atoi(t);

    if (stage == 2) {
        // reduce stage: loadIntermediateFile (main.cu:441) + reduce (:447-473)
        FILE *g = fopen(inter_path, "rb");
        if (!g) { perror(inter_path); return 1; }
        std::string text;
        char tmp[1 << 16];
        size_t got;
        while ((got = fread(tmp, 1, sizeof tmp, g)) > 0) text.append(tmp, got);
        fclose(g);
        locust_ctx *ctx = nullptr;
        int rc = locust_create(&cfg, &ctx);
        if (rc) return die(ctx, "locust_create", rc);
        rc = locust_map_intermediate(ctx, text.data(), text.size());
        if (rc) return die(ctx, "locust_map_intermediate", rc);
        rc = locust_process(ctx);
        if (rc) return die(ctx, "locust_process", rc);
        rc = locust_reduce(ctx);
        if (rc) return die(ctx, "locust_reduce", rc);
        printf("GPU reduce %d nanoseconds \n", (int)((locust_stage_ms(ctx, LOCUST_STAGE_PROCESS) +
                                                      locust_stage_ms(ctx, LOCUST_STAGE_REDUCE)) * 1e6f));
        const char *text_out = nullptr;
        uint64_t nb = 0;
        rc = locust_format_result(ctx, key_format, &text_out, &nb);
        if (rc) return die(ctx, "locust_format_result", rc);
        fwrite(text_out, 1, nb, stdout);
        locust_destroy(ctx);
        printf("\nDone\n");
        return 0;
    }

    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 1; }
    if (start_line < 0 && !getenv("LOCUST_DEVICES") && !getenv("LOCUST_WHOLE_FILE")) {
        // loadFile's role (main.cu:242-259) without holding the file: it is read in chunks into a ring of three pinned
        // buffers, each chunk cut after its last delimiter, and while chunk k is read from disk chunk k - 1 is on its
        // way to the GPU and being tokenised (locust_map copies and maps asynchronously).
        const size_t CH = getenv("LOCUST_CHUNK_MIB") ? (size_t)atol(getenv("LOCUST_CHUNK_MIB")) << 20 : (size_t)64 << 20;
        uint8_t *ring[3] = {nullptr, nullptr, nullptr};
        for (auto &b : ring)
            if (cudaMallocHost(reinterpret_cast<void **>(&b), CH) != cudaSuccess) {
                fprintf(stderr, "MapReduce: no CUDA device / pinned allocation failed (this build has no CPU path)\n");
                return 1;
            }
        locust_ctx *ctx = nullptr;
        int rc = locust_create(&cfg, &ctx);
        if (rc) return die(ctx, "locust_create", rc);
        bool isd[256] = {};
        for (const char *q = " ,.-;:'()\"\t\n"; *q; ++q) isd[(uint8_t)*q] = true;
        long nlines = 0;
        size_t carry = 0;
        bool last_byte_nl = true, any = false;
        for (int k = 0;; ++k) {
            uint8_t *b = ring[k % 3];
            const size_t got = fread(b + carry, 1, CH - carry, f);
            const size_t total = carry + got;
            const bool eof = got < CH - carry;
            size_t end = total;
            if (!eof) {
                while (end > 0 && !isd[b[end - 1]]) --end;
                if (end == 0) { fprintf(stderr, "MapReduce: a token longer than the read chunk; set LOCUST_WHOLE_FILE=1\n"); return 1; }
            }
            if (end && memchr(b, 0, end)) {
                fprintf(stderr, "MapReduce: input contains NUL bytes; the reference stops a line at NUL, this build refuses\n");
                return 1;
            }
            for (const uint8_t *q = b, *e = b + end; q < e;) {
                const void *nl = memchr(q, '\n', (size_t)(e - q));
                if (!nl) break;
                ++nlines;
                q = (const uint8_t *)nl + 1;
            }
            if (end) { last_byte_nl = b[end - 1] == '\n'; any = true; }
            rc = locust_wait_input(ctx);                 // chunk k - 1 has left its buffer: buffer (k + 1) % 3 is free
            if (rc) return die(ctx, "locust_wait_input", rc);
            if (end) { rc = locust_map(ctx, b, end, LOCUST_HOST); if (rc) return die(ctx, "locust_map", rc); }
            carry = total - end;
            if (carry) memcpy(ring[(k + 1) % 3], b + end, carry);
            if (eof) break;
        }
        fclose(f);
        if (any && !last_byte_nl) ++nlines;              // getline counts a last line without a newline
        printf("Length: %i\n", (int)nlines);
        rc = locust_process(ctx);
        if (rc) return die(ctx, "locust_process", rc);
        rc = locust_reduce(ctx);
        if (rc) return die(ctx, "locust_reduce", rc);
        printf("GPU mapping %d nanoseconds \n", (int)(locust_stage_ms(ctx, LOCUST_STAGE_MAP) * 1e6f));
        printf("GPU stream compaction and sorting %d nanoseconds \n", (int)(locust_stage_ms(ctx, LOCUST_STAGE_PROCESS) * 1e6f));
        const char *text = nullptr;
        uint64_t nbytes = 0;
        if (stage == 1) {
            rc = locust_format_result(ctx, LOCUST_FORMAT_INTERMEDIATE, &text, &nbytes);
            if (rc) return die(ctx, "locust_format_result", rc);
            FILE *g = fopen(inter_path, "wb");
            if (!g) { perror(inter_path); return 1; }
            fwrite(text, 1, nbytes, g);
            fclose(g);
            locust_destroy(ctx);
            printf("MODE_MULTI: Finished map\n");
            return 0;
        }
        printf("GPU reduce %d nanoseconds \n", (int)(locust_stage_ms(ctx, LOCUST_STAGE_REDUCE) * 1e6f));
        rc = locust_format_result(ctx, key_format, &text, &nbytes);
        if (rc) return die(ctx, "locust_format_result", rc);
        fwrite(text, 1, nbytes, stdout);
        locust_destroy(ctx);
        for (auto &b : ring) cudaFreeHost(b);
        printf("\nDone\n");
        return 0;
    }
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

    if (const char *devs = getenv("LOCUST_DEVICES")) {
        if (stage != 0) { fprintf(stderr, "MapReduce: LOCUST_DEVICES applies to single mode only\n"); return 2; }
        std::vector<int> devices;
        for (const char *q = devs; *q;) {
            char *end = nullptr;
            const long v = strtol(q, &end, 10);
            if (end == q) break;
            devices.push_back((int)v);
            q = (*end == ',') ? end + 1 : end;
        }
        if (devices.empty()) { fprintf(stderr, "MapReduce: LOCUST_DEVICES holds no device number\n"); return 2; }
        locust_multi *m = nullptr;
        int rcm = locust_multi_create(devices.data(), (int)devices.size(), &cfg, &m);
        if (rcm == LOCUST_OK) rcm = locust_multi_wordcount(m, buf + lo, hi - lo);
        if (rcm != LOCUST_OK) {
            fprintf(stderr, "MapReduce: multi-GPU job failed (%d): %s\n", rcm, m ? locust_multi_last_error(m) : "");
            if (m) locust_multi_destroy(m);
            return 1;
        }
        // stage times: the slowest rank's, as a launcher waiting for every node would see them
        float t_map = 0.f, t_proc = 0.f, t_red = 0.f;
        for (int r = 0; r < (int)devices.size(); ++r) {
            locust_ctx *cr = locust_multi_context(m, r);
            const float a = locust_stage_ms(cr, LOCUST_STAGE_MAP), b = locust_stage_ms(cr, LOCUST_STAGE_PROCESS), c = locust_stage_ms(cr, LOCUST_STAGE_REDUCE);
            if (a > t_map) t_map = a;
            if (b > t_proc) t_proc = b;
            if (c > t_red) t_red = c;
        }
        printf("GPU mapping %d nanoseconds \n", (int)(t_map * 1e6f));
        printf("GPU stream compaction and sorting %d nanoseconds \n", (int)(t_proc * 1e6f));
        printf("GPU reduce %d nanoseconds \n", (int)(t_red * 1e6f));
        const char *text = nullptr;
        uint64_t nbytes = 0;
        locust_ctx *c0 = locust_multi_context(m, 0);
        rcm = locust_format_result(c0, key_format, &text, &nbytes);
        if (rcm) { fprintf(stderr, "MapReduce: locust_format_result failed (%d): %s\n", rcm, locust_last_error(c0)); return 1; }
        fwrite(text, 1, nbytes, stdout);
        locust_multi_destroy(m);
        cudaFreeHost(buf);
        printf("\nDone\n");
        return 0;
    }

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
    if (stage == 1) {
        // map stage of multi mode: partials to the intermediate file, then exit (main.cu:421-432)
        const char *lines = nullptr;
        uint64_t nb = 0;
        rc = locust_format_result(ctx, LOCUST_FORMAT_INTERMEDIATE, &lines, &nb);
        if (rc) return die(ctx, "locust_format_result", rc);
        FILE *g = fopen(inter_path, "wb");
        if (!g) { perror(inter_path); return 1; }
        fwrite(lines, 1, nb, g);
        fclose(g);
        locust_destroy(ctx);
        cudaFreeHost(buf);
        printf("MODE_MULTI: Finished map\n");
        return 0;
    }
    printf("GPU reduce %d nanoseconds \n", (int)(locust_stage_ms(ctx, LOCUST_STAGE_REDUCE) * 1e6f));

    const char *text = nullptr;
    uint64_t nbytes = 0;
    rc = locust_format_result(ctx, key_format, &text, &nbytes);
    if (rc) return die(ctx, "locust_format_result", rc);
    fwrite(text, 1, nbytes, stdout);
    locust_destroy(ctx);
    cudaFreeHost(buf);
    printf("\nDone\n");
    return 0;
}
