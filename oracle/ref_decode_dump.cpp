// ref_decode_dump -- golden-vector driver over the UNMODIFIED reference's public llama.h (TEST INFRASTRUCTURE).
//
// Our own ~150 lines; linked against the reference objects built by oracle/Makefile into oracle/_ref/.  Single rank
// (the mini-MPI shim acts as a 1-rank world).  Replays a script of llama.h calls -- multi-sequence llama_decode
// batches and llama_kv_cache_seq_rm/cp/keep -- and dumps what the reference produced:
//   <out>.logits.bin   float32 [n_tokens][n_vocab] for every decode, concatenated   (llama_get_logits_ith)
//   <out>.kv.txt       one line per "KV" command: used_cells, then "cell:pos:seq,seq,..." for every non-empty cell
//                      (llama_kv_cache_view_update, /root/reference/llama.cpp:9160-9222), plus "D rc" lines.
//
// usage: ref_decode_dump model.gguf n_ctx max_seq script.txt out_prefix [n_threads]
// script:   D <n>            followed by n lines:  <token> <pos> <n_seq> <seq>...
//           RM <seq> <p0> <p1> | CP <src> <dst> <p0> <p1> | KEEP <seq> | KV
#include "llama.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

int main(int argc, char ** argv) {
    if (argc < 6) {
        fprintf(stderr, "usage: %s model.gguf n_ctx max_seq script.txt out_prefix [n_threads]\n", argv[0]);
        return 2;
    }
    const char * model_path = argv[1];
    const int    n_ctx      = atoi(argv[2]);
    const int    max_seq    = atoi(argv[3]);
    const char * script     = argv[4];
    std::string  out        = argv[5];
    const int    n_threads  = argc > 6 ? atoi(argv[6]) : 4;

    llama_backend_init(false);
    llama_model_params mp = llama_model_default_params();
    llama_model * model   = llama_load_model_from_file(model_path, mp);
    if (!model) return 1;
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx           = n_ctx;
    cp.n_batch         = 512;
    cp.n_threads       = n_threads;
    cp.n_threads_batch = n_threads;
    cp.logits_all      = true;
    llama_context * ctx = llama_new_context_with_model(model, cp);
    if (!ctx) return 1;
    const int n_vocab = llama_n_vocab(model);

    FILE * fs = fopen(script, "r");
    FILE * fl = fopen((out + ".logits.bin").c_str(), "wb");
    FILE * fk = fopen((out + ".kv.txt").c_str(), "w");
    if (!fs || !fl || !fk) return 1;

    llama_kv_cache_view view = llama_kv_cache_view_init(ctx, max_seq);
    llama_batch batch        = llama_batch_init(512, 0, max_seq);
    int  batch_id = 0;
    char op[16];
    while (fscanf(fs, "%15s", op) == 1) {
        if (!strcmp(op, "D")) {
            int n;
            if (fscanf(fs, "%d", &n) != 1) return 1;
            batch.n_tokens = n;
            for (int i = 0; i < n; i++) {
                int tok, pos, ns;
                if (fscanf(fs, "%d %d %d", &tok, &pos, &ns) != 3) return 1;
                batch.token[i]    = tok;
                batch.pos[i]      = pos;
                batch.n_seq_id[i] = ns;
                for (int s = 0; s < ns; s++) if (fscanf(fs, "%d", &batch.seq_id[i][s]) != 1) return 1;
                batch.logits[i] = 1;
            }
            batch.batch_id  = batch_id++;
            batch.max_n_seq = max_seq;
            const int rc = llama_decode(ctx, batch);
            fprintf(fk, "D %d\n", rc);
            if (rc == 0) {
                for (int i = 0; i < n; i++) fwrite(llama_get_logits_ith(ctx, i), sizeof(float), n_vocab, fl);
            }
        } else if (!strcmp(op, "RM")) {
            int s, p0, p1;
            if (fscanf(fs, "%d %d %d", &s, &p0, &p1) != 3) return 1;
            llama_kv_cache_seq_rm(ctx, s, p0, p1);
        } else if (!strcmp(op, "CP")) {
            int a, b, p0, p1;
            if (fscanf(fs, "%d %d %d %d", &a, &b, &p0, &p1) != 4) return 1;
            llama_kv_cache_seq_cp(ctx, a, b, p0, p1);
        } else if (!strcmp(op, "KEEP")) {
            int s;
            if (fscanf(fs, "%d", &s) != 1) return 1;
            llama_kv_cache_seq_keep(ctx, s);
        } else if (!strcmp(op, "KV")) {
            llama_kv_cache_view_update(ctx, &view);
            fprintf(fk, "KV %d", view.used_cells);
            for (int i = 0; i < view.n_cells; i++) {
                const llama_seq_id * cs = view.cells_sequences + (size_t) i * view.n_max_seq;
                if (cs[0] < 0) continue;
                fprintf(fk, " %d:%d:", i, view.cells[i].pos);
                for (int s = 0; s < view.n_max_seq && cs[s] >= 0; s++) fprintf(fk, s ? ",%d" : "%d", cs[s]);
            }
            fprintf(fk, "\n");
        } else {
            fprintf(stderr, "bad op %s\n", op);
            return 1;
        }
    }
    fclose(fs); fclose(fl); fclose(fk);
    fflush(stdout);
    _Exit(0);   // skip llama_free: the reference's ggml_mpi_free touches a freed context on a 1-rank world
}
