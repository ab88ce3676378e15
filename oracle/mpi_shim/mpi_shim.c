/*
 * mpi_shim.c -- mini-MPI over POSIX shared memory (TEST INFRASTRUCTURE, not product code).
 *
 * See mpi.h for why this exists.  One process per rank, started by mpirun_shim (same directory),
 * which creates a tmpfs file, initialises the header below and exports
 *   MPISHIM_SHM, MPISHIM_RANK, MPISHIM_SIZE.
 * Without those variables the library behaves as a 1-rank world (no shared memory at all), which is
 * how the single-process golden-vector drivers use the reference's llama.cpp.
 *
 * Every message is a node in the receiver's inbox list (arrival order == send order per sender, so
 * MPI's non-overtaking rule holds).  All state changes happen under one process-shared mutex; each
 * receiver sleeps on its own condition variable.  Collectives are built from point-to-point
 * messages that carry an "internal" flag so that user-level ANY_TAG probes never see them.
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#define MPISHIM_MAXR  32
#define MPISHIM_MAGIC 0x4d50495348494d31ull

struct mpishim_comm {
    int id;
    int rank;
    int size;
    int split_seq;
    int coll_seq;
    int ranks[MPISHIM_MAXR];   /* comm rank -> world rank */
};

struct mpishim_comm mpishim_world = { 1, 0, 1, 0, 0, { 0 } };

typedef struct {
    uint64_t        magic;
    pthread_mutex_t lock;
    pthread_cond_t  cond[MPISHIM_MAXR];
    int             size;
    uint64_t        arena_off;
    uint64_t        arena_size;
    uint64_t        free_head;
    uint64_t        inbox_head[MPISHIM_MAXR];
    uint64_t        inbox_tail[MPISHIM_MAXR];
    uint64_t        bytes_sent[MPISHIM_MAXR];
    uint64_t        msgs_sent[MPISHIM_MAXR];
} shm_hdr;

typedef struct {
    uint64_t size;   /* bytes, including this header */
    uint64_t next;   /* offset of next free block, 0 = none */
} free_blk;

typedef struct {
    uint64_t blk_size;
    uint64_t next;
    int32_t  comm_id;
    int32_t  src;      /* rank of the sender inside the communicator */
    int32_t  tag;
    int32_t  internal;
    int32_t  nbytes;
    int32_t  pad;
    char     payload[];
} msg_t;

static shm_hdr * g_hdr  = NULL;
static char *    g_base = NULL;
static int       g_rank = 0;
static int       g_size = 1;

#define AT(off) ((void *)(g_base + (off)))

size_t mpishim_header_bytes(void) { return (sizeof(shm_hdr) + 4095) & ~(size_t)4095; }

/* called by the launcher on a fresh mapping */
void mpishim_init_segment(void * base, size_t total, int nranks) {
    shm_hdr * h = (shm_hdr *) base;
    memset(h, 0, sizeof(*h));
    pthread_mutexattr_t ma;
    pthread_mutexattr_init(&ma);
    pthread_mutexattr_setpshared(&ma, PTHREAD_PROCESS_SHARED);
    pthread_mutex_init(&h->lock, &ma);
    pthread_condattr_t ca;
    pthread_condattr_init(&ca);
    pthread_condattr_setpshared(&ca, PTHREAD_PROCESS_SHARED);
    for (int i = 0; i < MPISHIM_MAXR; i++) pthread_cond_init(&h->cond[i], &ca);
    h->size       = nranks;
    h->arena_off  = mpishim_header_bytes();
    h->arena_size = total - h->arena_off;
    free_blk * fb = (free_blk *)((char *) base + h->arena_off);
    fb->size      = h->arena_size;
    fb->next      = 0;
    h->free_head  = h->arena_off;
    h->magic      = MPISHIM_MAGIC;
}

/* ---- first-fit allocator with address-ordered free list; caller holds the lock ---- */
static uint64_t arena_alloc(uint64_t need) {
    need = (need + 63) & ~63ull;
    uint64_t prev = 0, cur = g_hdr->free_head;
    while (cur) {
        free_blk * b = (free_blk *) AT(cur);
        if (b->size >= need) {
            if (b->size - need >= 128) {
                uint64_t rest = cur + need;
                free_blk * r  = (free_blk *) AT(rest);
                r->size = b->size - need;
                r->next = b->next;
                if (prev) ((free_blk *) AT(prev))->next = rest; else g_hdr->free_head = rest;
                b->size = need;
            } else {
                if (prev) ((free_blk *) AT(prev))->next = b->next; else g_hdr->free_head = b->next;
            }
            return cur;
        }
        prev = cur;
        cur  = b->next;
    }
    return 0;
}

static void arena_free(uint64_t off, uint64_t size) {
    uint64_t prev = 0, cur = g_hdr->free_head;
    while (cur && cur < off) { prev = cur; cur = ((free_blk *) AT(cur))->next; }
    free_blk * b = (free_blk *) AT(off);
    b->size = size;
    b->next = cur;
    if (prev) ((free_blk *) AT(prev))->next = off; else g_hdr->free_head = off;
    if (cur && off + b->size == cur) {           /* merge with the following block */
        free_blk * n = (free_blk *) AT(cur);
        b->size += n->size;
        b->next  = n->next;
    }
    if (prev) {                                  /* merge with the preceding block */
        free_blk * p = (free_blk *) AT(prev);
        if (prev + p->size == off) { p->size += b->size; p->next = b->next; }
    }
}

/* ---- point to point ---- */
static void put_msg(MPI_Comm comm, int dst, int tag, int internal, const void * buf, int nbytes) {
    if (!g_hdr) { fprintf(stderr, "mpishim: send in a 1-rank world ignored (dst=%d tag=%d)\n", dst, tag); return; }
    const int wdst = comm->ranks[dst];
    pthread_mutex_lock(&g_hdr->lock);
    uint64_t off;
    int spins = 0;
    while ((off = arena_alloc(sizeof(msg_t) + (uint64_t) nbytes)) == 0) {
        pthread_mutex_unlock(&g_hdr->lock);
        if (++spins > 60000) { fprintf(stderr, "mpishim: arena exhausted\n"); abort(); }
        usleep(1000);
        pthread_mutex_lock(&g_hdr->lock);
    }
    msg_t * m   = (msg_t *) AT(off);
    uint64_t bs = ((free_blk *) m)->size;
    m->blk_size = bs;
    m->next     = 0;
    m->comm_id  = comm->id;
    m->src      = comm->rank;
    m->tag      = tag;
    m->internal = internal;
    m->nbytes   = nbytes;
    if (nbytes) memcpy(m->payload, buf, (size_t) nbytes);
    if (g_hdr->inbox_tail[wdst]) ((msg_t *) AT(g_hdr->inbox_tail[wdst]))->next = off;
    else g_hdr->inbox_head[wdst] = off;
    g_hdr->inbox_tail[wdst] = off;
    g_hdr->bytes_sent[g_rank] += (uint64_t) nbytes;
    g_hdr->msgs_sent[g_rank]  += 1;
    pthread_cond_broadcast(&g_hdr->cond[wdst]);
    pthread_mutex_unlock(&g_hdr->lock);
}

/* find the first matching message in my inbox; caller holds the lock. *pprev receives its predecessor */
static uint64_t find_msg(MPI_Comm comm, int src, int tag, int internal, uint64_t * pprev) {
    uint64_t prev = 0, cur = g_hdr->inbox_head[g_rank];
    while (cur) {
        msg_t * m = (msg_t *) AT(cur);
        if (m->comm_id == comm->id && m->internal == internal &&
            (src == MPI_ANY_SOURCE || m->src == src) && (tag == MPI_ANY_TAG || m->tag == tag)) {
            *pprev = prev;
            return cur;
        }
        prev = cur;
        cur  = m->next;
    }
    return 0;
}

static void get_msg(MPI_Comm comm, int src, int tag, int internal, void * buf, int maxbytes, MPI_Status * st, int remove) {
    if (!g_hdr) { fprintf(stderr, "mpishim: recv in a 1-rank world ignored (src=%d tag=%d)\n", src, tag); return; }
    pthread_mutex_lock(&g_hdr->lock);
    uint64_t prev, off;
    while ((off = find_msg(comm, src, tag, internal, &prev)) == 0) {
        pthread_cond_wait(&g_hdr->cond[g_rank], &g_hdr->lock);
    }
    msg_t * m = (msg_t *) AT(off);
    if (st) { st->MPI_SOURCE = m->src; st->MPI_TAG = m->tag; st->MPI_ERROR = 0; st->mpishim_bytes = m->nbytes; }
    if (remove) {
        int n = m->nbytes < maxbytes ? m->nbytes : maxbytes;
        if (n) memcpy(buf, m->payload, (size_t) n);
        if (prev) ((msg_t *) AT(prev))->next = m->next; else g_hdr->inbox_head[g_rank] = m->next;
        if (g_hdr->inbox_tail[g_rank] == off) g_hdr->inbox_tail[g_rank] = prev;
        arena_free(off, m->blk_size);
    }
    pthread_mutex_unlock(&g_hdr->lock);
}

/* ---- API ---- */
int MPI_Init_thread(int * argc, char *** argv, int required, int * provided) {
    (void) argc; (void) argv;
    if (provided) *provided = required;
    if (g_hdr) return MPI_SUCCESS;
    const char * shm = getenv("MPISHIM_SHM");
    const char * rk  = getenv("MPISHIM_RANK");
    const char * sz  = getenv("MPISHIM_SIZE");
    if (!shm || !rk || !sz || atoi(sz) <= 1) {
        g_rank = 0; g_size = 1;
        mpishim_world.rank = 0; mpishim_world.size = 1; mpishim_world.ranks[0] = 0;
        return MPI_SUCCESS;
    }
    g_rank = atoi(rk);
    g_size = atoi(sz);
    int fd = open(shm, O_RDWR);
    if (fd < 0) { perror("mpishim: open"); abort(); }
    struct stat sb;
    fstat(fd, &sb);
    g_base = (char *) mmap(NULL, (size_t) sb.st_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (g_base == MAP_FAILED) { perror("mpishim: mmap"); abort(); }
    close(fd);
    g_hdr = (shm_hdr *) g_base;
    if (g_hdr->magic != MPISHIM_MAGIC || g_hdr->size != g_size) { fprintf(stderr, "mpishim: bad segment\n"); abort(); }
    mpishim_world.rank = g_rank;
    mpishim_world.size = g_size;
    for (int i = 0; i < g_size; i++) mpishim_world.ranks[i] = i;
    return MPI_SUCCESS;
}

int MPI_Finalize(void) { return MPI_SUCCESS; }

int MPI_Comm_rank(MPI_Comm comm, int * rank) { *rank = comm->rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm comm, int * size) { *size = comm->size; return MPI_SUCCESS; }

int MPI_Buffer_attach(void * buf, int size)   { (void) buf; (void) size; return MPI_SUCCESS; }
int MPI_Buffer_detach(void * buf, int * size) { (void) buf; (void) size; return MPI_SUCCESS; }

int MPI_Bsend(const void * buf, int count, MPI_Datatype dt, int dst, int tag, MPI_Comm comm) {
    put_msg(comm, dst, tag, 0, buf, count * dt);
    return MPI_SUCCESS;
}

int MPI_Recv(void * buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Status * st) {
    get_msg(comm, src, tag, 0, buf, count * dt, st, 1);
    return MPI_SUCCESS;
}

int MPI_Probe(int src, int tag, MPI_Comm comm, MPI_Status * st) {
    get_msg(comm, src, tag, 0, NULL, 0, st, 0);
    return MPI_SUCCESS;
}

int MPI_Iprobe(int src, int tag, MPI_Comm comm, int * flag, MPI_Status * st) {
    *flag = 0;
    if (!g_hdr) return MPI_SUCCESS;
    pthread_mutex_lock(&g_hdr->lock);
    uint64_t prev, off = find_msg(comm, src, tag, 0, &prev);
    if (off) {
        msg_t * m = (msg_t *) AT(off);
        *flag = 1;
        if (st) { st->MPI_SOURCE = m->src; st->MPI_TAG = m->tag; st->MPI_ERROR = 0; st->mpishim_bytes = m->nbytes; }
    }
    pthread_mutex_unlock(&g_hdr->lock);
    return MPI_SUCCESS;
}

int MPI_Get_count(const MPI_Status * st, MPI_Datatype dt, int * count) {
    *count = st->mpishim_bytes / dt;
    return MPI_SUCCESS;
}

int MPI_Irecv(void * buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Request * req) {
    req->buf = buf; req->count = count; req->dt = dt; req->src = src; req->tag = tag; req->comm = comm; req->active = 1;
    return MPI_SUCCESS;
}

int MPI_Wait(MPI_Request * req, MPI_Status * st) {
    if (req->active) {
        MPI_Recv(req->buf, req->count, req->dt, req->src, req->tag, req->comm, st);
        req->active = 0;
    }
    return MPI_SUCCESS;
}

/* collectives: internal point-to-point messages, tag = per-communicator collective sequence number */
int MPI_Barrier(MPI_Comm comm) {
    const int t = comm->coll_seq++;
    if (comm->size <= 1) return MPI_SUCCESS;
    char c = 0;
    if (comm->rank == 0) {
        for (int i = 1; i < comm->size; i++) get_msg(comm, i, t, 1, &c, 1, NULL, 1);
        for (int i = 1; i < comm->size; i++) put_msg(comm, i, t, 1, &c, 1);
    } else {
        put_msg(comm, 0, t, 1, &c, 1);
        get_msg(comm, 0, t, 1, &c, 1, NULL, 1);
    }
    return MPI_SUCCESS;
}

int MPI_Bcast(void * buf, int count, MPI_Datatype dt, int root, MPI_Comm comm) {
    const int t = comm->coll_seq++;
    if (comm->size <= 1) return MPI_SUCCESS;
    if (comm->rank == root) {
        for (int i = 0; i < comm->size; i++) if (i != root) put_msg(comm, i, t, 1, buf, count * dt);
    } else {
        get_msg(comm, root, t, 1, buf, count * dt, NULL, 1);
    }
    return MPI_SUCCESS;
}

int MPI_Scatter(const void * sbuf, int scount, MPI_Datatype sdt, void * rbuf, int rcount, MPI_Datatype rdt,
                int root, MPI_Comm comm) {
    const int t = comm->coll_seq++;
    if (comm->rank == root) {
        for (int i = 0; i < comm->size; i++) {
            const char * p = (const char *) sbuf + (size_t) i * scount * sdt;
            if (i == root) memcpy(rbuf, p, (size_t) rcount * rdt);
            else put_msg(comm, i, t, 1, p, scount * sdt);
        }
    } else {
        get_msg(comm, root, t, 1, rbuf, rcount * rdt, NULL, 1);
    }
    return MPI_SUCCESS;
}

int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm * newcomm) {
    const int seq = comm->split_seq++;
    const int t   = comm->coll_seq++;
    int table[MPISHIM_MAXR][2];
    if (comm->size <= 1) {
        table[0][0] = color; table[0][1] = key;
    } else if (comm->rank == 0) {
        table[0][0] = color; table[0][1] = key;
        for (int i = 1; i < comm->size; i++) get_msg(comm, i, t, 1, table[i], 8, NULL, 1);
        for (int i = 1; i < comm->size; i++) put_msg(comm, i, t, 1, table, (int) sizeof(int) * 2 * comm->size);
    } else {
        int mine[2] = { color, key };
        put_msg(comm, 0, t, 1, mine, 8);
        get_msg(comm, 0, t, 1, table, (int) sizeof(int) * 2 * comm->size, NULL, 1);
    }
    if (color == MPI_UNDEFINED) { *newcomm = MPI_COMM_NULL; return MPI_SUCCESS; }
    struct mpishim_comm * nc = (struct mpishim_comm *) calloc(1, sizeof(*nc));
    int members[MPISHIM_MAXR], n = 0;
    for (int i = 0; i < comm->size; i++) if (table[i][0] == color) members[n++] = i;
    for (int a = 1; a < n; a++) {                 /* stable insertion sort by key */
        int m = members[a], b = a - 1;
        while (b >= 0 && table[members[b]][1] > table[m][1]) { members[b + 1] = members[b]; b--; }
        members[b + 1] = m;
    }
    nc->size = n;
    nc->rank = -1;
    for (int i = 0; i < n; i++) {
        nc->ranks[i] = comm->ranks[members[i]];
        if (members[i] == comm->rank) nc->rank = i;
    }
    nc->id = (int)(((unsigned) comm->id * 1000003u + (unsigned)(seq + 1) * 7919u + (unsigned)(color + 17) * 104729u) & 0x7fffffffu);
    if (nc->id <= 1) nc->id += 2;
    *newcomm = nc;
    return MPI_SUCCESS;
}

int MPI_Comm_free(MPI_Comm * comm) {
    if (*comm && *comm != MPI_COMM_WORLD) free(*comm);
    *comm = MPI_COMM_NULL;
    return MPI_SUCCESS;
}

/* not MPI: traffic counters for the NVLink-bytes-per-hop comparison (bytes this rank has sent) */
unsigned long long mpishim_bytes_sent(void) { return g_hdr ? (unsigned long long) g_hdr->bytes_sent[g_rank] : 0ull; }
unsigned long long mpishim_msgs_sent(void)  { return g_hdr ? (unsigned long long) g_hdr->msgs_sent[g_rank]  : 0ull; }
