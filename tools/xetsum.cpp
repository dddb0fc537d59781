// xetsum — print or check Xet file hashes on a B200, like sha256sum.  A plain C++ consumer of the
// C ABI (include/xetgpu.h), i.e. what OME's Go wrappers do through cgo.
//
//   xetsum FILE...            print "<64-hex xet hash>  <path>" per file
//   xetsum -c LISTFILE        verify lines of that form; exit 1 on any mismatch
//   xetsum --chunks FILE      additionally print every chunk: "<end offset> <chunk hash>"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../include/xetgpu.h"

static int fail(XetError* e) {
    fprintf(stderr, "xetsum: error %d: %s%s%s\n", e->code, e->message ? e->message : "", e->details ? " — " : "",
            e->details ? e->details : "");
    int code = e->code;
    xetgpu_free_error(e);
    return code == XETGPU_ERR_CHECKSUM_MISMATCH ? 1 : 2;
}

int main(int argc, char** argv) {
    bool check = false, chunks = false;
    std::vector<const char*> paths;
    const char* list = nullptr;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "-c") && i + 1 < argc) {
            check = true;
            list = argv[++i];
        } else if (!strcmp(argv[i], "--chunks")) {
            chunks = true;
        } else {
            paths.push_back(argv[i]);
        }
    }
    if (!check && paths.empty()) {
        fprintf(stderr, "usage: xetsum [--chunks] FILE... | xetsum -c LISTFILE\n");
        return 2;
    }
    xetgpu_version_0_2_0();
    XetGpuContext* ctx = nullptr;
    XetGpuConfig cfg = {};
    cfg.device = -1;
    if (XetError* e = xetgpu_context_new(&cfg, &ctx)) return fail(e);
    int rc = 0;
    if (check) {
        FILE* f = fopen(list, "r");
        if (!f) {
            perror(list);
            return 2;
        }
        char line[8192];
        while (fgets(line, sizeof line, f)) {
            size_t n = strlen(line);
            while (n && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = 0;
            if (n < 67 || line[64] != ' ') continue;
            std::string hex(line, 64);
            const char* path = line + 66;
            XetError* e = xetgpu_verify_file(ctx, path, hex.c_str());
            if (!e) {
                printf("%s: OK\n", path);
            } else if (e->code == XETGPU_ERR_CHECKSUM_MISMATCH) {
                printf("%s: FAILED\n", path);
                xetgpu_free_error(e);
                rc = 1;
            } else {
                rc = fail(e);
            }
        }
        fclose(f);
    } else {
        XetGpuResult* res = nullptr;
        if (XetError* e = xetgpu_hash_files(ctx, paths.data(), paths.size(), chunks ? XETGPU_WANT_CHUNKS : 0, nullptr, &res))
            return fail(e);
        char hex[65];
        for (size_t i = 0; i < res->n_files; i++) {
            xetgpu_hash_to_hex(res->files[i].file_hash, hex);
            printf("%s  %s\n", hex, paths[i]);
            if (chunks)
                for (uint64_t c = 0; c < res->files[i].n_chunks; c++) {
                    uint64_t k = res->files[i].first_chunk + c;
                    xetgpu_hash_to_hex(res->chunk_hashes + 32 * k, hex);
                    printf("  %llu %s\n", (unsigned long long)res->chunk_ends[k], hex);
                }
        }
        xetgpu_free_result(res);
    }
    xetgpu_context_free(ctx);
    return rc;
}
