// pipeline.cu -- pipeline stages across GPUs: the replacement of the reference's ggml-mpi.c transport.
//
// One process per GPU.  Rank 0 is the head: it runs the scheduler and the draft model, owns the first layers of the
// target, and tells the other stages what to do.  Two channels replace the reference's 21 MPI tags (ggml-mpi.h:15-55):
//
//  * control ring -- a single-producer / multi-consumer byte ring in POSIX shared memory (one node).  The head appends
//    one record per decode (batch descriptors + the resolved KV op-log), per finished run, per cancel, per logit-bias
//    change; every stage consumes all records in order.  This gives the property the reference gets from per-sender MPI
//    ordering (ggml-mpi.c:173-234 sync_pipelined): every stage observes every KV op and decode in issue order, so the
//    KV metadata replicas stay identical.  The stages' HOST code needs these records to enqueue kernels; nothing in
//    them is read by a GPU of another process, so they do not travel over NVLink.
//  * data plane -- NCCL send/recv over NVLink, enqueued in the compute stream: activations [n_tokens][n_embd] fp32 to the
//    next stage (reference ggml_mpi_tensor_send / _recv, ggml-mpi.c:451-503), and the sampled rows (top-8 ids / logits /
//    probabilities, 96 B per row) from the last stage back to the head (the reference returns n_vocab floats per row,
//    llama.cpp:5757-5786).  Two communicators so that the head's result receives (own stream) never queue behind its
//    activation sends.
//
// Cancellation (reference GGML_MPI_CANCEL_RUN, llama.cpp:9981-9993): the record reaches every stage's host through the
// ring and the stage writes its device cancel table from its high-priority stream; kernels of a cancelled run exit at
// their first instruction, sends and receives still take place (payload ignored) so the pipeline stays in step.
#include "engine.cuh"

#include <dlfcn.h>
#include <fcntl.h>
#include <nccl.h>
#include <sched.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <vector>

// ---------------------------------------------------------------------------------------------------------------------
// control ring
#define PIPE_MAX_RANKS 16
#define PIPE_MAGIC     0x31424950u
enum { REC_PAD = 0, REC_DECODE = 1, REC_FINISH = 2, REC_CANCEL = 3, REC_BIAS = 4, REC_RETURN = 5, REC_SHUTDOWN = 6, REC_INTS = 7, REC_LOGITS = 8 };

struct alignas(64) ring_word { std::atomic<uint64_t> v; char pad[56]; };
struct ring_hdr {
    ring_word head;                      // bytes published by the head rank (monotonic)
    ring_word tail[PIPE_MAX_RANKS];      // bytes consumed by each stage
    ring_word bar_count, bar_gen;        // sense-reversing barrier over all ranks
    uint64_t  cap;
    uint32_t  n_ranks;
    std::atomic<uint32_t> magic;
    uint8_t   nccl_ids[256];             // the two ncclUniqueIds, published by rank 0 (launchers without a store of their own)
    std::atomic<uint32_t> ids_ready;
    uint64_t  created_us;                // wall clock of the creation: lets a stage tell a fresh ring from one a crashed run left behind
    // peer-store hop: every stage publishes the IPC handle of its hop buffer here (generation = how many pipelines' worth of
    // contexts this ring has seen, so that a handle of an earlier context is never taken for the current one)
    cudaIpcMemHandle_t    hop_ipc[PIPE_MAX_RANKS];
    std::atomic<uint32_t> hop_gen[PIPE_MAX_RANKS];
    uint32_t              hop_embd[PIPE_MAX_RANKS];
};
struct rec_hdr { uint32_t bytes, kind; };
struct rec_decode { int32_t batch_id, n_tokens, n_chunks, n_ops; uint32_t desc_stride, pad; };

struct pipe_state {
    int        rank = 0, size = 1;
    ring_hdr * h = nullptr;
    uint8_t *  data = nullptr;
    size_t     map_bytes = 0;
    std::string name;
    uint64_t   wpos = 0;                 // producer cursor
    uint32_t   open_bytes = 0;           // record being written
    // NCCL
    void *     nccl_lib = nullptr;
    ncclComm_t fwd = nullptr, back = nullptr;
    cudaStream_t res_stream = nullptr;
    uint64_t   hop_bytes = 0, hop_msgs = 0, back_bytes = 0;
    // last int record seen by the ggml_mpi_* shim
    int        last_tag = -1, last_count = 0;
};
static pipe_state P;

static double now_s() { struct timeval tv; gettimeofday(&tv, nullptr); return tv.tv_sec + 1e-6 * tv.tv_usec; }

// spin, then yield; gives up after `PIB_PIPE_TIMEOUT` seconds (default 600) so a dead peer cannot hang a GPU box for ever
struct spinner {
    int    n = 0;
    double t0 = 0;
    bool   step() {
        if (++n < 2000) { __builtin_ia32_pause(); return true; }
        sched_yield();
        if ((n & 0xfff) == 0) {
            if (t0 == 0) t0 = now_s();
            static double limit = getenv("PIB_PIPE_TIMEOUT") ? atof(getenv("PIB_PIPE_TIMEOUT")) : 600.0;
            if (now_s() - t0 > limit) return false;
        }
        return true;
    }
};

static uint64_t g_ring_min_created_us = 0;      // pib_pipe_open_env: refuse rings older than this (same name = same MASTER_PORT)

extern "C" int pib_pipe_ring_open(const char * name, int rank, int n_ranks, size_t cap_bytes) {
    if (P.h) { pib_set_error("pipe: ring already open"); return PIB_ERR_ARG; }
    if (n_ranks < 1 || n_ranks > PIPE_MAX_RANKS || rank < 0 || rank >= n_ranks) { pib_set_error("pipe: bad rank %d / %d", rank, n_ranks); return PIB_ERR_ARG; }
    cap_bytes = (cap_bytes + 4095) & ~(size_t) 4095;
    const size_t total = 4096 + cap_bytes;
    static_assert(sizeof(ring_hdr) <= 4096, "ring header must fit its page");
    int fd = -1;
    if (rank == 0) {
        shm_unlink(name);
        fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd < 0 || ftruncate(fd, (off_t) total) != 0) { pib_set_error("pipe: cannot create shared ring %s", name); if (fd >= 0) close(fd); return PIB_ERR_NOMEM; }
    } else {
        spinner sp;
        while (true) {
            fd = shm_open(name, O_RDWR, 0600);
            struct stat st;
            if (fd >= 0 && fstat(fd, &st) == 0 && (size_t) st.st_size >= total) break;
            if (fd >= 0) close(fd);
            usleep(1000);
            if (!sp.step()) { pib_set_error("pipe: shared ring %s never appeared", name); return PIB_ERR_ARG; }
        }
    }
    void * p = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) { pib_set_error("pipe: mmap of %s failed", name); return PIB_ERR_NOMEM; }
    P.h = (ring_hdr *) p; P.data = (uint8_t *) p + 4096; P.map_bytes = total; P.rank = rank; P.size = n_ranks; P.name = name; P.wpos = 0;
    if (rank == 0) {
        P.h->cap = cap_bytes; P.h->n_ranks = (uint32_t) n_ranks;
        P.h->head.v.store(0); P.h->bar_count.v.store(0); P.h->bar_gen.v.store(0); P.h->ids_ready.store(0);
        P.h->created_us = (uint64_t) (now_s() * 1e6);
        for (int i = 0; i < PIPE_MAX_RANKS; i++) { P.h->tail[i].v.store(0); P.h->hop_gen[i].store(0); }
        P.h->magic.store(PIPE_MAGIC, std::memory_order_release);
    } else {
        spinner sp;
        while (P.h->magic.load(std::memory_order_acquire) != PIPE_MAGIC) if (!sp.step()) { pib_set_error("pipe: ring %s never initialised", name); return PIB_ERR_ARG; }
        if (P.h->created_us < g_ring_min_created_us) {
            // a segment of this name left behind by a run that died: the head is about to unlink and re-create it
            munmap((void *) P.h, P.map_bytes);
            P.h = nullptr; P.data = nullptr;
            usleep(20000);
            static int retries = 0;
            if (++retries > 3000) { pib_set_error("pipe: only a stale ring %s is there", name); return PIB_ERR_ARG; }
            return pib_pipe_ring_open(name, rank, n_ranks, cap_bytes);
        }
    }
    return PIB_OK;
}

extern "C" int pib_pipe_barrier(void) {
    if (!P.h) return PIB_OK;
    const uint64_t gen = P.h->bar_gen.v.load(std::memory_order_acquire);
    if (P.h->bar_count.v.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint64_t) P.size) {
        P.h->bar_count.v.store(0, std::memory_order_relaxed);
        P.h->bar_gen.v.store(gen + 1, std::memory_order_release);
    } else {
        spinner sp;
        while (P.h->bar_gen.v.load(std::memory_order_acquire) == gen) if (!sp.step()) { pib_set_error("pipe: barrier timed out"); return PIB_ERR_ARG; }
    }
    return PIB_OK;
}

extern "C" void pib_pipe_ring_close(void) {
    if (!P.h) return;
    pib_pipe_barrier();
    const bool owner = P.rank == 0;
    munmap((void *) P.h, P.map_bytes);
    if (owner) shm_unlink(P.name.c_str());
    P.h = nullptr; P.data = nullptr; P.size = 1; P.rank = 0;
}

extern "C" int pib_pipe_rank(void) { return P.rank; }
extern "C" int pib_pipe_size(void) { return P.size; }

static bool wait_space(uint64_t need) {
    spinner sp;
    while (true) {
        uint64_t mn = UINT64_MAX;
        for (int r = 1; r < P.size; r++) { const uint64_t t = P.h->tail[r].v.load(std::memory_order_acquire); if (t < mn) mn = t; }
        if (P.size == 1 || P.wpos + need - mn <= P.h->cap) return true;
        if (!sp.step()) return false;
    }
}

// producer: reserve a record of `payload` bytes; returns the payload pointer (nullptr: ring closed or consumers stuck)
static uint8_t * rec_begin(uint32_t kind, size_t payload) {
    if (!P.h || P.rank != 0) return nullptr;
    const uint64_t need = (sizeof(rec_hdr) + payload + 7) & ~(uint64_t) 7;
    if (need > P.h->cap / 2) { pib_set_error("pipe: record of %zu bytes does not fit the control ring", payload); return nullptr; }
    uint64_t pos = P.wpos % P.h->cap;
    if (P.h->cap - pos < need) {
        const uint64_t pad = P.h->cap - pos;
        if (!wait_space(pad)) { pib_set_error("pipe: control ring full (a stage stopped consuming)"); return nullptr; }
        rec_hdr * ph = (rec_hdr *) (P.data + pos);
        ph->bytes = (uint32_t) pad; ph->kind = REC_PAD;
        P.wpos += pad;
        P.h->head.v.store(P.wpos, std::memory_order_release);
        pos = 0;
    }
    if (!wait_space(need)) { pib_set_error("pipe: control ring full (a stage stopped consuming)"); return nullptr; }
    rec_hdr * h = (rec_hdr *) (P.data + pos);
    h->bytes = (uint32_t) need; h->kind = kind;
    P.open_bytes = (uint32_t) need;
    return (uint8_t *) (h + 1);
}
static void rec_commit() {
    P.wpos += P.open_bytes;
    P.open_bytes = 0;
    P.h->head.v.store(P.wpos, std::memory_order_release);
}

// consumer: next record (blocking unless `nowait`); PAD records are skipped
static const rec_hdr * rec_peek(bool nowait) {
    if (!P.h || P.rank == 0) return nullptr;
    spinner sp;
    while (true) {
        const uint64_t t = P.h->tail[P.rank].v.load(std::memory_order_relaxed);
        if (P.h->head.v.load(std::memory_order_acquire) == t) {
            if (nowait) return nullptr;
            if (!sp.step()) { pib_set_error("pipe: no control record for too long (head gone?)"); return nullptr; }
            continue;
        }
        const rec_hdr * h = (const rec_hdr *) (P.data + t % P.h->cap);
        if (h->kind == REC_PAD) { P.h->tail[P.rank].v.store(t + h->bytes, std::memory_order_release); continue; }
        return h;
    }
}
static void rec_pop(const rec_hdr * h) {
    const uint64_t t = P.h->tail[P.rank].v.load(std::memory_order_relaxed);
    P.h->tail[P.rank].v.store(t + h->bytes, std::memory_order_release);
}

// ---- records written by the head (called from engine.cu) -------------------------------------------------------------
bool pib_pipe_ring_active() { return P.h != nullptr && P.size > 1; }

int pib_pipe_ship_decode(int batch_id, int n_tokens, int n_chunks, const uint8_t * const * descs, size_t desc_stride,
                         const int * slots, const KvOp * ops, int n_ops) {
    if (!pib_pipe_ring_active() || P.rank != 0) return PIB_OK;
    const size_t payload = sizeof(rec_decode) + (size_t) n_chunks * desc_stride + (size_t) n_ops * sizeof(KvOp);
    uint8_t * p = rec_begin(REC_DECODE, payload);
    if (!p) return PIB_ERR_NOMEM;
    rec_decode * d = (rec_decode *) p;
    d->batch_id = batch_id; d->n_tokens = n_tokens; d->n_chunks = n_chunks; d->n_ops = n_ops; d->desc_stride = (uint32_t) desc_stride; d->pad = 0;
    uint8_t * q = p + sizeof(rec_decode);
    for (int ch = 0; ch < n_chunks; ch++, q += desc_stride) memcpy(q, descs[ch], desc_stride);
    KvOp * o = (KvOp *) q;
    for (int i = 0; i < n_ops; i++) {
        o[i] = ops[i];
        if (o[i].op == KVOP_ALLOC) {                 // descriptor slot of the head -> chunk index inside this record
            int ch = 0;
            while (ch < n_chunks && slots[ch] != ops[i].p1) ch++;
            o[i].p1 = ch;
        }
    }
    rec_commit();
    return PIB_OK;
}

int pib_pipe_ship_ints(int kind, int tag, const int32_t * vals, int n) {
    if (!pib_pipe_ring_active() || P.rank != 0) return PIB_OK;
    uint8_t * p = rec_begin((uint32_t) kind, sizeof(int32_t) * (size_t) (n + 2));
    if (!p) return PIB_ERR_NOMEM;
    int32_t * q = (int32_t *) p;
    q[0] = tag; q[1] = n;
    if (n) memcpy(q + 2, vals, sizeof(int32_t) * (size_t) n);
    rec_commit();
    return PIB_OK;
}

// cancel record: the batch ids, then the head's monotonic clock (ns) at the moment of the decision, for the latency statistics below
static int64_t mono_ns() { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (int64_t) ts.tv_sec * 1000000000ll + ts.tv_nsec; }
int pib_pipe_ship_cancel(const int32_t * batch_ids, int n) {
    if (!pib_pipe_ring_active() || P.rank != 0) return PIB_OK;
    uint8_t * p = rec_begin(REC_CANCEL, sizeof(int32_t) * (size_t) (n + 4));
    if (!p) return PIB_ERR_NOMEM;
    int32_t * q = (int32_t *) p;
    q[0] = 0; q[1] = n;
    memcpy(q + 2, batch_ids, sizeof(int32_t) * (size_t) n);
    const int64_t t = mono_ns();
    memcpy(q + 2 + n, &t, 8);
    rec_commit();
    return PIB_OK;
}
// stage side: head decision -> record seen by this stage's host -> cancel word written on this stage's GPU (the latter only with
// PIB_CANCEL_TIMING=1: it synchronises the control stream after every cancel)
static struct { uint64_t n = 0; double host_sum = 0, host_max = 0, dev_sum = 0, dev_max = 0; uint64_t n_dev = 0; } g_cancel_stats;
extern "C" void pib_pipe_cancel_stats(uint64_t * n, double * host_mean_us, double * host_max_us, double * dev_mean_us, double * dev_max_us) {
    *n = g_cancel_stats.n;
    *host_mean_us = g_cancel_stats.n ? g_cancel_stats.host_sum / g_cancel_stats.n : 0; *host_max_us = g_cancel_stats.host_max;
    *dev_mean_us = g_cancel_stats.n_dev ? g_cancel_stats.dev_sum / g_cancel_stats.n_dev : 0; *dev_max_us = g_cancel_stats.dev_max;
}

int pib_pipe_ship_bias(const int32_t * ids, const float * vals, int n) {
    if (!pib_pipe_ring_active() || P.rank != 0) return PIB_OK;
    uint8_t * p = rec_begin(REC_BIAS, sizeof(int32_t) * (size_t) (2 * n + 2));
    if (!p) return PIB_ERR_NOMEM;
    int32_t * q = (int32_t *) p;
    q[0] = 0; q[1] = n;
    memcpy(q + 2, ids, sizeof(int32_t) * (size_t) n);
    memcpy(q + 2 + n, vals, sizeof(float) * (size_t) n);
    rec_commit();
    return PIB_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// NCCL, resolved at run time (the process may already have torch's copy of libnccl.so.2 loaded; dlopen returns that one)
struct nccl_fns {
    ncclResult_t (*GetUniqueId)(ncclUniqueId *);
    ncclResult_t (*CommInitRankConfig)(ncclComm_t *, int, ncclUniqueId, int, ncclConfig_t *);
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)(void);
    ncclResult_t (*GroupEnd)(void);
    const char * (*GetErrorString)(ncclResult_t);
};
static nccl_fns N;

static int nccl_load() {
    if (P.nccl_lib) return PIB_OK;
    setenv("NCCL_DEBUG_FILE", "/dev/stderr", 0);      // NCCL's banner / debug lines must not land in a CLI's stdout (generated text only)
    const char * path = getenv("PIB_NCCL_LIB");
    void * lib = dlopen(path ? path : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { pib_set_error("pipe: cannot load NCCL (%s); set PIB_NCCL_LIB", dlerror()); return PIB_ERR_UNSUPP; }
#define SYM(field, name) do { *(void **) &N.field = dlsym(lib, name); if (!N.field) { pib_set_error("pipe: NCCL symbol %s missing", name); return PIB_ERR_UNSUPP; } } while (0)
    SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommDestroy, "ncclCommDestroy");
    SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv"); SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    *(void **) &N.CommInitRankConfig = dlsym(lib, "ncclCommInitRankConfig");
    P.nccl_lib = lib;
    return PIB_OK;
}
#define NCCL_CHECK(expr) do { ncclResult_t r__ = (expr); if (r__ != ncclSuccess) { pib_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, N.GetErrorString(r__)); return PIB_ERR_CUDA; } } while (0)

extern "C" int pib_pipe_unique_id(uint8_t * out128) {
    if (nccl_load() != PIB_OK) return PIB_ERR_UNSUPP;
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    NCCL_CHECK(N.GetUniqueId(&id));
    memcpy(out128, &id, 128);
    return PIB_OK;
}

// both communicators span all ranks; call after pib_init on every rank, collectively
extern "C" int pib_pipe_nccl_init(int rank, int size, const uint8_t * id_fwd, const uint8_t * id_back) {
    if (nccl_load() != PIB_OK) return PIB_ERR_UNSUPP;
    if (P.h && (P.rank != rank || P.size != size)) { pib_set_error("pipe: NCCL rank/size differ from the control ring's"); return PIB_ERR_ARG; }
    P.rank = rank; P.size = size;
    ncclUniqueId a, b;
    memcpy(&a, id_fwd, 128); memcpy(&b, id_back, 128);
    // one CTA per transfer: the messages are small (<= 512 KiB) and a receive that waits for the pipeline should not hold
    // more than one SM slot of the head GPU
    if (N.CommInitRankConfig) {
        ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
        cfg.minCTAs = 1; cfg.maxCTAs = 1;
        NCCL_CHECK(N.CommInitRankConfig(&P.fwd, size, a, rank, &cfg));
        ncclConfig_t cfg2 = NCCL_CONFIG_INITIALIZER;
        cfg2.minCTAs = 1; cfg2.maxCTAs = 1;
        NCCL_CHECK(N.CommInitRankConfig(&P.back, size, b, rank, &cfg2));
    } else {
        NCCL_CHECK(N.CommInitRank(&P.fwd, size, a, rank));
        NCCL_CHECK(N.CommInitRank(&P.back, size, b, rank));
    }
    int lo, hi;
    PIB_CHECK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    PIB_CHECK(cudaStreamCreateWithPriority(&P.res_stream, cudaStreamNonBlocking, hi));
    return PIB_OK;
}

// Everything from the launcher's environment (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_PORT as torchrun / mpirun-style
// launchers set them): control ring named after the port, NCCL ids handed over through the ring's header page.  For
// programs that have no rendezvous store of their own (the `speculative` CLI); returns the world size, < 0 on error.
extern "C" int pib_pipe_open_env(void) {
    const char * r = getenv("RANK"), * w = getenv("WORLD_SIZE"), * lr = getenv("LOCAL_RANK"), * port = getenv("MASTER_PORT");
    const int rank = r ? atoi(r) : 0, world = w ? atoi(w) : 1;
    if (world <= 1) return 1;
    if (pib_init(lr ? atoi(lr) : rank) != PIB_OK) return PIB_ERR_CUDA;
    const std::string name = std::string("/pib_ring_") + (port ? port : "29500");
    g_ring_min_created_us = (uint64_t) (now_s() * 1e6) - 120ull * 1000000ull;       // the ranks of one launch start within two minutes
    int rc = pib_pipe_ring_open(name.c_str(), rank, world, (size_t) 64 << 20);
    if (rc != PIB_OK) return rc;
    if (rank == 0) {
        if ((rc = pib_pipe_unique_id(P.h->nccl_ids)) != PIB_OK || (rc = pib_pipe_unique_id(P.h->nccl_ids + 128)) != PIB_OK) return rc;
        P.h->ids_ready.store(1, std::memory_order_release);
    } else {
        spinner sp;
        while (P.h->ids_ready.load(std::memory_order_acquire) != 1) if (!sp.step()) { pib_set_error("pipe: NCCL ids never published"); return PIB_ERR_ARG; }
    }
    uint8_t ids[256];
    memcpy(ids, P.h->nccl_ids, 256);
    if ((rc = pib_pipe_nccl_init(rank, world, ids, ids + 128)) != PIB_OK) return rc;
    return world;
}

extern "C" void pib_pipe_nccl_free(void) {
    if (P.fwd) N.CommDestroy(P.fwd);
    if (P.back) N.CommDestroy(P.back);
    if (P.res_stream) cudaStreamDestroy(P.res_stream);
    P.fwd = P.back = nullptr; P.res_stream = nullptr;
}

extern "C" void pib_pipe_traffic(uint64_t * hop_bytes, uint64_t * hop_msgs, uint64_t * back_bytes) {
    if (hop_bytes) *hop_bytes = P.hop_bytes;
    if (hop_msgs) *hop_msgs = P.hop_msgs;
    if (back_bytes) *back_bytes = P.back_bytes;
}

bool pib_pipe_nccl_active() { return P.fwd != nullptr && P.size > 1; }

// ---------------------------------------------------------------------------------------------------------------------
// peer-store hop (reference: ggml_mpi_tensor_send / _recv of the layer output, ggml-mpi.c:451-503, 710-721).  Every stage but the
// first owns a ring of HOP_SLOTS activation slots + one flag per slot in ITS OWN memory, exported with cudaIpc and mapped by the
// stage in front of it.  The last kernel of a stage forms the stage's output (residual add) and stores it straight into the next
// stage's slot with ordinary st.global over NVLink, then releases the slot's flag (system scope); the first kernel of the next
// stage acquires the flag, copies the slot into its work buffer and acknowledges into the SENDER's memory (back pressure: the
// sender never runs more than HOP_SLOTS chunks ahead).  Both kernels take everything that varies from device memory (sequence
// counters), so they live inside the stage's captured CUDA graph: no ncclSend / ncclRecv launch pair per stage.  Plain loads and
// stores only.
#define HOP_SLOTS 32
struct hop_dev {                          // first 256 bytes of a stage's exported allocation
    unsigned int flags[HOP_SLOTS];        // written by the PREVIOUS stage: 1 + sequence number of the chunk in the slot
    unsigned int ack;                     // written by the NEXT stage: chunks it has consumed
    unsigned int send_seq, recv_seq;      // this stage's own counters (chunks sent on / received)
    unsigned int pad[64 - HOP_SLOTS - 3];
};
static_assert(sizeof(hop_dev) == 256, "hop header");

__device__ __forceinline__ unsigned int hop_ld_acquire(const unsigned int * p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void hop_st_release(unsigned int * p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// out = a [+ b [+ c3]] written into the next stage's slot; one CTA (the payload is 32 KiB per token)
__global__ void __launch_bounds__(1024) hop_send_kernel(const float4 * __restrict__ a, const float4 * __restrict__ b, const float4 * __restrict__ c3,
                                                        int n4, hop_dev * mine, hop_dev * peer, size_t slot_f4) {
    pib_pdl_wait();
    pib_pdl_trigger();
    const unsigned int seq = mine->send_seq, slot = seq % HOP_SLOTS;
    if (threadIdx.x == 0) while ((int) (seq - hop_ld_acquire(&mine->ack)) >= HOP_SLOTS) { }
    __syncthreads();
    float4 * dst = reinterpret_cast<float4 *>(reinterpret_cast<uint8_t *>(peer) + 256) + (size_t) slot * slot_f4;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 v = a[i];
        if (b) { const float4 w = b[i]; v.x = __fadd_rn(v.x, w.x); v.y = __fadd_rn(v.y, w.y); v.z = __fadd_rn(v.z, w.z); v.w = __fadd_rn(v.w, w.w); }
        if (c3) { const float4 w = c3[i]; v.x = __fadd_rn(v.x, w.x); v.y = __fadd_rn(v.y, w.y); v.z = __fadd_rn(v.z, w.z); v.w = __fadd_rn(v.w, w.w); }
        dst[i] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) { hop_st_release(&peer->flags[slot], seq + 1); mine->send_seq = seq + 1; }
}

__global__ void __launch_bounds__(1024) hop_recv_kernel(float4 * __restrict__ x, int n4, hop_dev * mine, hop_dev * peer_prev, size_t slot_f4) {
    pib_pdl_wait();
    pib_pdl_trigger();
    const unsigned int seq = mine->recv_seq, slot = seq % HOP_SLOTS;
    if (threadIdx.x == 0) while (hop_ld_acquire(&mine->flags[slot]) != seq + 1) { }
    __syncthreads();
    const float4 * src = reinterpret_cast<const float4 *>(reinterpret_cast<const uint8_t *>(mine) + 256) + (size_t) slot * slot_f4;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) x[i] = __ldcv(src + i);       // the slot was written from outside this GPU: not through L1
    __syncthreads();
    if (threadIdx.x == 0) { mine->recv_seq = seq + 1; hop_st_release(&peer_prev->ack, seq + 1); }
}

static uint32_t g_hop_generation = 0;

// collective over the stages of the pipeline (each creates its target context at the same point of the program)
int pib_pipe_hop_setup(pib_context * c) {
    const char * mode = getenv("PIB_HOP");
    if (!P.h || c->stage.size <= 1 || (mode && !strcmp(mode, "nccl"))) return PIB_OK;
    const int      r = c->stage.rank, n = c->stage.size;
    const uint32_t gen = ++g_hop_generation;
    const size_t   slot_bytes = (size_t) PIB_CHUNK * c->model->n_embd * sizeof(float);
    const size_t   bytes = 256 + (r > 0 ? (size_t) HOP_SLOTS * slot_bytes : 0);
    int dev = 0;
    PIB_CHECK(cudaGetDevice(&dev));
    PIB_CHECK(cudaMalloc(&c->hop_mine, bytes));
    PIB_CHECK(cudaMemset(c->hop_mine, 0, bytes));
    PIB_CHECK(cudaDeviceSynchronize());
    PIB_CHECK(cudaIpcGetMemHandle(&P.h->hop_ipc[r], c->hop_mine));
    P.h->hop_embd[r] = (uint32_t) c->model->n_embd;
    P.h->hop_gen[r].store(gen, std::memory_order_release);
    auto open_peer = [&](int peer, void ** out) -> int {
        spinner sp;
        while (P.h->hop_gen[peer].load(std::memory_order_acquire) != gen) if (!sp.step()) { pib_set_error("pipe: stage %d never published its hop buffer", peer); return PIB_ERR_ARG; }
        if (P.h->hop_embd[peer] != (uint32_t) c->model->n_embd) { pib_set_error("pipe: stage %d runs another model width", peer); return PIB_ERR_ARG; }
        cudaIpcMemHandle_t hnd = P.h->hop_ipc[peer];
        const cudaError_t e = cudaIpcOpenMemHandle(out, hnd, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { cudaGetLastError(); *out = nullptr; pib_set_error("pipe: cannot map stage %d's hop buffer: %s", peer, cudaGetErrorString(e)); return PIB_ERR_CUDA; }
        return PIB_OK;
    };
    int rc = PIB_OK;
    if (r + 1 < n) rc = open_peer(r + 1, &c->hop_next);
    if (rc == PIB_OK && r > 0) rc = open_peer(r - 1, &c->hop_prev);
    c->hop_slot_f4 = slot_bytes / 16;
    // every stage reports through the ring header (a stage that failed publishes generation | 0x80000000): if any mapping failed
    // -- no peer access between two of the GPUs -- ALL stages fall back to the NCCL hop
    P.h->hop_gen[r].store(rc == PIB_OK ? gen : (gen | 0x80000000u), std::memory_order_release);
    pib_pipe_barrier();
    bool all_ok = true;
    for (int i = 0; i < n; i++) if (P.h->hop_gen[i].load(std::memory_order_acquire) != gen) all_ok = false;
    pib_pipe_barrier();
    c->hop_peer = all_ok;
    if (!all_ok && r == 0) fprintf(stderr, "pipe: peer mapping unavailable (%s); the hop uses ncclSend / ncclRecv\n", pib_last_error());
    return PIB_OK;
}

void pib_pipe_hop_free(pib_context * c) {
    if (c->hop_next) cudaIpcCloseMemHandle(c->hop_next);
    if (c->hop_prev) cudaIpcCloseMemHandle(c->hop_prev);
    if (c->hop_mine) cudaFree(c->hop_mine);
    c->hop_next = c->hop_prev = c->hop_mine = nullptr;
    c->hop_peer = false;
}

// first kernel of a stage > 0: the previous stage's output -> c->x
int pib_hop_recv_launch(pib_context * c, int n_tokens, cudaStream_t st) {
    const int n4 = n_tokens * c->model->n_embd / 4;
    pib_launch(hop_recv_kernel, dim3(1), dim3(1024), 0, st, reinterpret_cast<float4 *>(c->x), n4, (hop_dev *) c->hop_mine, (hop_dev *) c->hop_prev, c->hop_slot_f4);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}
// last kernel of a stage < last: a [+ b [+ c3]] -> the next stage's slot
int pib_hop_send_launch(pib_context * c, const float * a, const float * b, const float * c3, int n_tokens, cudaStream_t st) {
    const int n4 = n_tokens * c->model->n_embd / 4;
    pib_launch(hop_send_kernel, dim3(1), dim3(1024), 0, st, reinterpret_cast<const float4 *>(a), reinterpret_cast<const float4 *>(b),
               reinterpret_cast<const float4 *>(c3), n4, (hop_dev *) c->hop_mine, (hop_dev *) c->hop_next, c->hop_slot_f4);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}
// bytes this stage stores into the next stage's memory: counted where the chunk is enqueued (a graph replay does not pass through
// the launch function above)
void pib_pipe_hop_account(size_t bytes) { P.hop_bytes += bytes; P.hop_msgs++; }

// engine hooks ---------------------------------------------------------------------------------------------------------
static int hop_recv(pib_context * c, float * x, int n_tokens, const int * cancel) {
    (void) cancel;
    const size_t n = (size_t) n_tokens * c->model->n_embd;
    NCCL_CHECK(N.Recv(x, n, ncclFloat, P.rank - 1, P.fwd, c->stream));
    return PIB_OK;
}
static int hop_send(pib_context * c, const float * x, int n_tokens, const int * cancel) {
    (void) cancel;
    const size_t n = (size_t) n_tokens * c->model->n_embd;
    NCCL_CHECK(N.Send(x, n, ncclFloat, P.rank + 1, P.fwd, c->stream));
    P.hop_bytes += n * 4; P.hop_msgs++;
    return PIB_OK;
}
// what comes back from the last stage: the result of the device-side acceptance walk when the batch carried a table (80 bytes),
// else the sampled rows (96 bytes per row).  Both ends know which: the table travels in the descriptor.
int pib_pipe_send_results(pib_context * c, pib_run & run) {
    const size_t bytes = run.has_walk ? sizeof(pib_accept_result) : sizeof(pib_row_result) * (size_t) run.n_tokens;
    NCCL_CHECK(N.Send(run.has_walk ? (const void *) run.acc_dev : (const void *) run.res_dev, bytes, ncclUint8, 0, P.back, c->stream));
    P.back_bytes += bytes;
    return PIB_OK;
}
int pib_pipe_post_result_recv(pib_context * c, pib_run & run) {
    const size_t bytes = run.has_walk ? sizeof(pib_accept_result) : sizeof(pib_row_result) * (size_t) run.n_tokens;
    void * dev = run.has_walk ? (void *) run.acc_dev : (void *) run.res_dev, * host = run.has_walk ? (void *) run.acc_host : (void *) run.res_host;
    run.rows_on_host = !run.has_walk;
    NCCL_CHECK(N.Recv(dev, bytes, ncclUint8, P.size - 1, P.back, P.res_stream));
    PIB_CHECK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, P.res_stream));
    c->d2h_bytes += bytes;
    PIB_CHECK(cudaEventRecord(run.done, P.res_stream));
    return PIB_OK;
}
// head: one full logits row of the last finished run, from the last stage (llama_get_logits_ith for callers that still read
// n_vocab floats, e.g. the reference's common/sampling.cpp:131).  Ordered behind the result rows of the runs in flight.
int pib_pipe_fetch_logits(pib_context * c, int row, float * dev_staging, float * host_dst) {
    const int32_t r = row;
    int rc = pib_pipe_ship_ints(REC_LOGITS, 0, &r, 1);
    if (rc != PIB_OK) return rc;
    const size_t V = (size_t) c->model->n_vocab;
    NCCL_CHECK(N.Recv(dev_staging, V, ncclFloat, P.size - 1, P.back, P.res_stream));
    PIB_CHECK(cudaMemcpyAsync(host_dst, dev_staging, V * 4, cudaMemcpyDeviceToHost, P.res_stream));
    PIB_CHECK(cudaStreamSynchronize(P.res_stream));
    c->d2h_bytes += V * 4;
    return PIB_OK;
}
// last stage: answer it
static int send_logits_row(pib_context * c, int row) {
    if (!c->stage.last()) return PIB_OK;
    if (c->last_finished < 0) { pib_set_error("stage: logits requested before any run finished"); return PIB_ERR_ARG; }
    pib_run & r = c->runs[c->last_finished];
    if (row < 0 || row >= r.n_tokens || !r.logits) { pib_set_error("stage: bad logits row %d", row); return PIB_ERR_ARG; }
    const size_t V = (size_t) c->model->n_vocab;
    NCCL_CHECK(N.Send(r.logits + (size_t) row * V, V, ncclFloat, 0, P.back, c->stream));
    P.back_bytes += V * 4;
    return PIB_OK;
}

void pib_pipe_install_hooks() {
    pib_hop_recv = hop_recv;
    pib_hop_send = hop_send;
}

// ---------------------------------------------------------------------------------------------------------------------
// the stage loop of ranks > 0 (reference worker loop: llama.cpp:9941-9977 llama_process_mpi_worker + the transaction
// switch at :9944-9976).  Returns 0 when the head paused the pipeline (REC_RETURN; stream drained), 1 on shutdown.
int pib_stage_serve(pib_context * c) {
    if (!P.h || P.rank == 0) { pib_set_error("pipe: serve is for ranks > 0 of an open pipeline"); return -1; }
    while (true) {
        const rec_hdr * h = rec_peek(false);
        if (!h) return -1;
        const uint8_t * p = (const uint8_t *) (h + 1);
        int rc = PIB_OK, ret = -2;
        switch (h->kind) {
            case REC_DECODE: {
                const rec_decode * d = (const rec_decode *) p;
                const uint8_t * descs = p + sizeof(rec_decode);
                const KvOp *    ops   = (const KvOp *) (descs + (size_t) d->n_chunks * d->desc_stride);
                rc = pib_decode_replay(c, d->batch_id, d->n_tokens, d->n_chunks, descs, d->desc_stride, ops, d->n_ops);
            } break;
            case REC_FINISH: { const int32_t * q = (const int32_t *) p; for (int i = 0; i < q[1]; i++) pib_stage_finish(c, q[2 + i]); } break;
            case REC_CANCEL: {
                const int32_t * q = (const int32_t *) p;
                int64_t t0;
                memcpy(&t0, q + 2 + q[1], 8);
                const double host_us = (mono_ns() - t0) / 1000.0;
                g_cancel_stats.n++; g_cancel_stats.host_sum += host_us; g_cancel_stats.host_max = std::max(g_cancel_stats.host_max, host_us);
                pib_cancel(c, q + 2, q[1]);
                static const bool timing = getenv("PIB_CANCEL_TIMING") && atoi(getenv("PIB_CANCEL_TIMING")) != 0;
                if (timing) {
                    cudaStreamSynchronize(c->ctrl);
                    const double dev_us = (mono_ns() - t0) / 1000.0;
                    g_cancel_stats.n_dev++; g_cancel_stats.dev_sum += dev_us; g_cancel_stats.dev_max = std::max(g_cancel_stats.dev_max, dev_us);
                }
            } break;
            case REC_BIAS:   { const int32_t * q = (const int32_t *) p; pib_set_logit_bias(c, q + 2, (const float *) (q + 2 + q[1]), q[1]); } break;
            case REC_INTS:   { const int32_t * q = (const int32_t *) p; P.last_tag = q[0]; P.last_count = q[1]; } break;
            case REC_LOGITS: { const int32_t * q = (const int32_t *) p; rc = send_logits_row(c, q[2]); } break;
            case REC_RETURN:   ret = 0; break;
            case REC_SHUTDOWN: ret = 1; break;
            default: pib_set_error("pipe: unknown control record %u", h->kind); rc = PIB_ERR_ARG;
        }
        rec_pop(h);
        if (rc != PIB_OK) return -1;
        if (ret >= 0) {
            if (cudaStreamSynchronize(c->stream) != cudaSuccess) { pib_set_error("pipe: stage stream failed: %s", cudaGetErrorString(cudaGetLastError())); return -1; }
            return ret;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// CPU-only test hooks: the same records, applied to a host mirror instead of a GPU replica (tests/test_pipeline_cpu.py)
void pib_kvtest_ship_decode(pib_kvtest * t, int n_tokens, const int32_t * pos, const uint32_t * tokbits) {
    const int    W = t->kv.n_words, n_chunks = (n_tokens + PIB_CHUNK - 1) / PIB_CHUNK;
    const size_t stride = pib_desc_stride(W);
    std::vector<uint8_t>         buf((size_t) n_chunks * stride, 0);
    std::vector<const uint8_t *> dp(n_chunks);
    std::vector<int>             slots(n_chunks);
    const uint32_t head = t->kv.head;
    for (int ch = 0; ch < n_chunks; ch++) {
        const int t0 = ch * PIB_CHUNK, nt = std::min(PIB_CHUNK, n_tokens - t0);
        uint8_t * d  = buf.data() + (size_t) ch * stride;
        int32_t * hdr = reinterpret_cast<int32_t *>(d + PIB_DESC_HDR_OFF);
        hdr[PIB_DH_NTOK] = nt; hdr[PIB_DH_KVHEAD] = (int32_t) (head + t0); hdr[PIB_DH_NKV] = (int32_t) t->kv.n;
        memcpy(d + PIB_DESC_POS_OFF, pos + t0, sizeof(int32_t) * nt);
        memcpy(d + PIB_DESC_BITS_OFF, tokbits + (size_t) t0 * W, sizeof(uint32_t) * (size_t) nt * W);
        dp[ch] = d; slots[ch] = ch;
        t->devops.push_back({ KVOP_ALLOC, (int32_t) (head + t0), nt, (int32_t) t->kv.n, ch, 0 });
    }
    t->devops.push_back({ KVOP_ADVANCE, n_tokens, (int32_t) head, 0, 0, 0 });
    pib_pipe_ship_decode(0, n_tokens, n_chunks, dp.data(), stride, slots.data(), t->devops.data(), (int) t->devops.size());
    t->devops.clear();
}

extern "C" void pib_kvtest_ship(void * h, int enable) { ((pib_kvtest *) h)->ship = enable != 0; }
// head: publish whatever the op-log still holds (ops issued after the last decode), then let the stages return
extern "C" int  pib_kvtest_pause(void * hh) {
    pib_kvtest * t = (pib_kvtest *) hh;
    if (t) {
        kvlog_flush(t->kv, t->pending, t->devops);
        if (!t->devops.empty()) pib_pipe_ship_decode(0, 0, 0, nullptr, pib_desc_stride(t->kv.n_words), nullptr, t->devops.data(), (int) t->devops.size());
        t->devops.clear();
    }
    return pib_pipe_ship_ints(REC_RETURN, 0, nullptr, 0);
}
extern "C" int  pib_kvtest_replay(void * hh) {
    pib_kvtest * t = (pib_kvtest *) hh;
    while (true) {
        const rec_hdr * h = rec_peek(false);
        if (!h) return -1;
        const uint8_t * p = (const uint8_t *) (h + 1);
        int ret = -2;
        if (h->kind == REC_DECODE) {
            const rec_decode * d = (const rec_decode *) p;
            const uint8_t * descs = p + sizeof(rec_decode);
            const KvOp *    ops   = (const KvOp *) (descs + (size_t) d->n_chunks * d->desc_stride);
            for (int i = 0; i < d->n_ops; i++) {
                const uint8_t * dd = ops[i].op == KVOP_ALLOC ? descs + (size_t) ops[i].p1 * d->desc_stride : descs;
                t->kv.apply(ops[i], reinterpret_cast<const int32_t *>(dd + PIB_DESC_POS_OFF), reinterpret_cast<const uint32_t *>(dd + PIB_DESC_BITS_OFF));
            }
        } else if (h->kind == REC_RETURN) ret = 0;
        else if (h->kind == REC_SHUTDOWN) ret = 1;
        rec_pop(h);
        if (ret >= 0) return ret;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// layer partition: reference ggml_mpi_split_range (ggml-mpi.c:523-559), called with start = 0, end = n_layer - 1
// (llama.cpp:8922-8936): range_length = end - start + 1; ranges[i] = [next, min(end, next + w_i * range_length + start)]
// in float arithmetic truncated to uint16, next = ranges[i][1]; the last stage gets [next, end].  Stage i then runs layers
// [ranges[i][0], ranges[i][1]) and the last stage runs to the end of the graph (ggml-mpi.c:640-650).
extern "C" void pib_split_range(int n_stages, int start, int end, const float * weights, uint16_t * out_pairs) {
    const uint16_t range_length = (uint16_t) (end - start + 1);
    uint16_t       next_layer   = 0;
    for (int i = 0; i < n_stages - 1; i++) {
        out_pairs[2 * i] = next_layer;
        const float hi = (float) out_pairs[2 * i] + weights[i] * (float) range_length + (float) start;
        const uint16_t e = (uint16_t) ((float) end < hi ? (float) end : hi);
        out_pairs[2 * i + 1] = e;
        next_layer = e;
    }
    out_pairs[2 * (n_stages - 1)]     = next_layer;
    out_pairs[2 * (n_stages - 1) + 1] = (uint16_t) end;
}

// ---------------------------------------------------------------------------------------------------------------------
// ggml_mpi_* entry points kept for callers of the reference's transport API (ggml-mpi.h:70-301), mapped onto the ring and
// NCCL.  Tensors and graphs do not exist here; the graph hooks (graph_creation_post / graph_compute_pre / _post) are what
// run_chunk does with pib_hop_recv / pib_hop_send.
struct ggml_mpi_context { int rank, size; int trans_id, recv_trans_id; uint16_t layer_start, layer_end; };

extern "C" {
void ggml_mpi_backend_init(void) {}                                          // ggml-mpi.c:58 (MPI_Init_thread): the launcher did it
void ggml_mpi_backend_free(void) {}
struct ggml_mpi_context * ggml_mpi_init(void) {                              // ggml-mpi.c:67
    ggml_mpi_context * ctx = new ggml_mpi_context();
    ctx->rank = P.rank; ctx->size = P.size; ctx->trans_id = 0; ctx->recv_trans_id = 0; ctx->layer_start = 0; ctx->layer_end = 0;
    return ctx;
}
void   ggml_mpi_free(struct ggml_mpi_context * ctx) { delete ctx; }          // :100
// :84-98 MPI_Comm_split: the reference uses it to separate the target pipeline's ranks from the draft pipeline's.  Both
// pipelines' heads live in rank 0's process here, so a split keeps the stage set (colour >= 0) or yields a lone context
struct ggml_mpi_context * ggml_mpi_split_comm(struct ggml_mpi_context * ctx, int color, int key) {
    (void) key;
    ggml_mpi_context * n = new ggml_mpi_context(*ctx);
    if (color < 0) { n->rank = 0; n->size = 1; }
    return n;
}
int    ggml_mpi_rank(struct ggml_mpi_context * ctx) { return ctx->rank; }   // :128
size_t ggml_mpi_size(struct ggml_mpi_context * ctx) { return (size_t) ctx->size; }
int    ggml_mpi_next_node(struct ggml_mpi_context * ctx) { return (ctx->rank + 1) % ctx->size; }      // :164
int    ggml_mpi_prev_node(struct ggml_mpi_context * ctx) { const int t = ctx->rank - 1; return t >= 0 ? t : ctx->size - 1; }
void   ggml_mpi_barrier(struct ggml_mpi_context * ctx) { (void) ctx; pib_pipe_barrier(); }           // :136
int    ggml_mpi_trans_id(struct ggml_mpi_context * ctx) { return ctx->trans_id; }                     // :38-48
int    ggml_mpi_recv_trans_id(struct ggml_mpi_context * ctx) { return ctx->recv_trans_id; }
void   ggml_mpi_inc_trans_id(struct ggml_mpi_context * ctx) { ctx->trans_id++; }

// :349-373: the head publishes `count` ints under `tag`; every other stage receives them in order
void ggml_mpi_sync_ints_pipelined(struct ggml_mpi_context * ctx, int32_t * vals, int count, int tag) {
    if (ctx->size < 2) return;
    if (ctx->rank == 0) { pib_pipe_ship_ints(REC_INTS, tag, vals, count); return; }
    const rec_hdr * h = rec_peek(false);
    if (!h || h->kind != REC_INTS) return;
    const int32_t * q = (const int32_t *) (h + 1);
    if (q[0] == tag) memcpy(vals, q + 2, sizeof(int32_t) * (size_t) (q[1] < count ? q[1] : count));
    P.last_tag = q[0]; P.last_count = q[1];
    rec_pop(h);
}
// :140-162
int ggml_mpi_iprobe(struct ggml_mpi_context * ctx, int src, int tag) {
    (void) src;
    if (ctx->rank == 0) return 0;
    const rec_hdr * h = rec_peek(true);
    if (!h || h->kind != REC_INTS) return 0;
    const int32_t * q = (const int32_t *) (h + 1);
    P.last_tag = q[0]; P.last_count = q[1];
    return tag < 0 || q[0] == tag;
}
void ggml_mpi_probe(struct ggml_mpi_context * ctx, int src, int tag) {
    (void) src; (void) tag;
    if (ctx->rank == 0) return;
    const rec_hdr * h = rec_peek(false);
    if (h && h->kind == REC_INTS) { const int32_t * q = (const int32_t *) (h + 1); P.last_tag = q[0]; P.last_count = q[1]; }
}
int ggml_mpi_status_tag(struct ggml_mpi_context * ctx) { (void) ctx; return P.last_tag; }
int ggml_mpi_status_count_int32(struct ggml_mpi_context * ctx) { (void) ctx; return P.last_count; }

// :399-421 -- `val` is a DEVICE pointer here; the transfer is enqueued on the NCCL forward communicator's result stream
void ggml_mpi_send_float_array_async(struct ggml_mpi_context * ctx, float * val, int arr_size, int dest, int tag) {
    (void) ctx; (void) tag;
    if (!pib_pipe_nccl_active()) return;
    // the reference's transport aborts on a failed MPI call (MPI_ERRORS_ARE_FATAL); a silent loss here would hang the peer
    const ncclResult_t r = N.Send(val, (size_t) arr_size, ncclFloat, dest, P.back, P.res_stream);
    if (r != ncclSuccess) { fprintf(stderr, "ggml_mpi_send_float_array_async: %s\n", N.GetErrorString(r)); abort(); }
}
void ggml_mpi_recv_float_array(struct ggml_mpi_context * ctx, float * val, int arr_size, int src, int tag) {
    (void) ctx; (void) tag;
    if (!pib_pipe_nccl_active()) return;
    const ncclResult_t r = N.Recv(val, (size_t) arr_size, ncclFloat, src, P.back, P.res_stream);
    if (r != ncclSuccess) { fprintf(stderr, "ggml_mpi_recv_float_array: %s\n", N.GetErrorString(r)); abort(); }
    const cudaError_t e = cudaStreamSynchronize(P.res_stream);
    if (e != cudaSuccess) { fprintf(stderr, "ggml_mpi_recv_float_array: %s\n", cudaGetErrorString(e)); abort(); }
}

// :523-559; the outer array and the rows are malloc'd, the caller frees the outer array (llama.cpp:8934-8936)
uint16_t ** ggml_mpi_split_range(struct ggml_mpi_context * ctx, uint16_t start, uint16_t end, float node_weights[]) {
    if (ctx->rank != 0) return nullptr;                                      // only the head computes the split (:536-538)
    uint16_t ** ranges = (uint16_t **) malloc(sizeof(uint16_t *) * (size_t) ctx->size);
    std::vector<uint16_t> flat((size_t) ctx->size * 2);
    pib_split_range(ctx->size, start, end, node_weights, flat.data());
    for (int i = 0; i < ctx->size; i++) {
        ranges[i]    = (uint16_t *) malloc(sizeof(uint16_t) * 2);
        ranges[i][0] = flat[2 * i]; ranges[i][1] = flat[2 * i + 1];
    }
    return ranges;
}
// :561-589: every stage learns its own range (MPI_Scatter in the reference; here each rank indexes the same table)
void ggml_mpi_scatter_layers(struct ggml_mpi_context * ctx, uint16_t ** layer_ranges) {
    int32_t flat[2 * PIPE_MAX_RANKS] = { 0 };
    if (ctx->rank == 0 && layer_ranges) for (int i = 0; i < ctx->size; i++) { flat[2 * i] = layer_ranges[i][0]; flat[2 * i + 1] = layer_ranges[i][1]; }
    ggml_mpi_sync_ints_pipelined(ctx, flat, 2 * ctx->size, 7 /* GGML_MPI_TRANSFER_TENSORS slot reused for the layer table */);
    ctx->layer_start = (uint16_t) flat[2 * ctx->rank]; ctx->layer_end = (uint16_t) flat[2 * ctx->rank + 1];
}
void ggml_mpi_layer_range(struct ggml_mpi_context * ctx, uint16_t * start, uint16_t * end) { *start = ctx->layer_start; *end = ctx->layer_end; }

// :375-397: one value from `root` to everybody.  Only the head can publish on the ring, so root must be 0.
void ggml_mpi_synch_int(struct ggml_mpi_context * ctx, int32_t * val, int root) {
    if (ctx->size < 2) return;
    if (root != 0) { pib_set_error("ggml_mpi_synch_int: only root 0 (the head) can publish"); return; }
    ggml_mpi_sync_ints_pipelined(ctx, val, 1, 100);
}
void ggml_mpi_synch_float(struct ggml_mpi_context * ctx, float * val, int root) {
    int32_t bits;
    memcpy(&bits, val, 4);
    ggml_mpi_synch_int(ctx, &bits, root);
    memcpy(val, &bits, 4);
}
// :361-373: the reverse direction (last stage -> head) carries no host records here: sampled rows and logits come back over
// NCCL, KV ops originate on the head only.  Kept as a no-op so that callers link.
void ggml_mpi_sync_ints_pipelined_back(struct ggml_mpi_context * ctx, int32_t * vals, int count, int tag) { (void) ctx; (void) vals; (void) count; (void) tag; }
bool ggml_mpi_is_decoding(struct ggml_mpi_context * ctx) { (void) ctx; return false; }          // :50-56 (state of the reference's async recv)
void ggml_mpi_wait_recv(struct ggml_mpi_context * ctx) { (void) ctx; }                            // :516-521: receives complete in stream order

// :236-347: the head publishes a batch (batch id, n_tokens, logits mask, tokens, n_seq_id, positions, flattened sequence ids,
// in the reference's message order and with its tags); every other stage receives it into the caller's pre-allocated
// arrays (the reference makes the same assumption, :293-300).  On stages the logits mask replaces *logits as in :270-290.
bool ggml_mpi_eval_init(struct ggml_mpi_context * ctx, int32_t * n_tokens, int32_t ** tokens, int32_t ** pos, int32_t ** n_seq_ids,
                        int32_t *** seq_id, int8_t ** logits, int32_t * batch_id, bool receive_only) {
    (void) receive_only;
    if (ctx->size < 2) return true;
    ggml_mpi_sync_ints_pipelined(ctx, batch_id, 1, 12 /* GGML_MPI_BATCH_ID */);
    ggml_mpi_sync_ints_pipelined(ctx, n_tokens, 1, 13 /* GGML_MPI_N_TOKENS */);
    const int n = *n_tokens;
    std::vector<int32_t> lg((size_t) n, 0);
    if (ctx->rank == 0 && *logits) for (int i = 0; i < n; i++) lg[i] = (*logits)[i];
    ggml_mpi_sync_ints_pipelined(ctx, lg.data(), n, 20 /* GGML_MPI_BATCH_LOGITS */);
    if (ctx->rank != 0) {
        bool any = false;
        for (int i = 0; i < n; i++) any = any || lg[i];
        if (*logits) { free(*logits); *logits = nullptr; }
        if (any) { *logits = (int8_t *) calloc((size_t) n, 1); for (int i = 0; i < n; i++) (*logits)[i] = (int8_t) lg[i]; }
    }
    ggml_mpi_sync_ints_pipelined(ctx, *tokens, n, 14 /* GGML_MPI_TOKENS */);
    ggml_mpi_sync_ints_pipelined(ctx, *n_seq_ids, n, 15 /* GGML_MPI_N_SEQ_IDS */);
    int total = 0;
    for (int i = 0; i < n; i++) total += (*n_seq_ids)[i];
    std::vector<int32_t> flat((size_t) total, 0);
    if (ctx->rank == 0) { int k = 0; for (int i = 0; i < n; i++) for (int j = 0; j < (*n_seq_ids)[i]; j++) flat[k++] = (*seq_id)[i][j]; }
    ggml_mpi_sync_ints_pipelined(ctx, *pos, n, 17 /* GGML_MPI_POS */);
    ggml_mpi_sync_ints_pipelined(ctx, flat.data(), total, 16 /* GGML_MPI_SEQ_IDS */);
    { int k = 0; for (int i = 0; i < n; i++) for (int j = 0; j < (*n_seq_ids)[i]; j++) (*seq_id)[i][j] = flat[k++]; }
    return true;
}

// :591-720 graph hooks: there is no ggml graph here; run_chunk (engine.cu) receives before the first local layer and sends
// after the last one through pib_hop_recv / pib_hop_send.  Present so that callers link.
void ggml_mpi_graph_creation_post(struct ggml_mpi_context * ctx, struct ggml_cgraph * gf, int n_layers) { (void) ctx; (void) gf; (void) n_layers; }
bool ggml_mpi_graph_compute_pre(struct ggml_mpi_context * ctx, struct ggml_cgraph * gf) { (void) ctx; (void) gf; return true; }
void ggml_mpi_graph_compute_post(struct ggml_mpi_context * ctx, struct ggml_cgraph * gf) { (void) ctx; (void) gf; }
}
