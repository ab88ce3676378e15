/*
 * mpi.h -- mini-MPI shim (TEST INFRASTRUCTURE, not product code).
 *
 * The reference's pipeline transport (/root/reference/ggml-mpi.c:5) includes <mpi.h>, but this
 * image ships no MPI and nothing can be installed.  This header + mpi_shim.c provide exactly the
 * 18 entry points and the constants ggml-mpi.c uses (enumerated in SURVEY.md §8c), implemented as
 * a single-host, multi-process message queue in POSIX shared memory.  Semantics honoured:
 *   - per-(sender, communicator) non-overtaking order (what the worker loop relies on,
 *     /root/reference/llama.cpp:9941-9977),
 *   - MPI_Bsend never blocks (the reference attaches a 128 MiB buffer, ggml-mpi.c:77-79),
 *   - MPI_Probe / MPI_Iprobe with ANY_SOURCE / ANY_TAG and MPI_Get_count.
 * It is only ever linked into binaries built under oracle/_ref/.
 */
#ifndef MPISHIM_MPI_H
#define MPISHIM_MPI_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mpishim_comm * MPI_Comm;
typedef int MPI_Datatype;           /* value == element size in bytes */

typedef struct {
    int MPI_SOURCE;
    int MPI_TAG;
    int MPI_ERROR;
    int mpishim_bytes;
} MPI_Status;

typedef struct {
    void *       buf;
    int          count;
    MPI_Datatype dt;
    int          src;
    int          tag;
    MPI_Comm     comm;
    int          active;
} MPI_Request;

extern struct mpishim_comm mpishim_world;

#define MPI_COMM_WORLD      (&mpishim_world)
#define MPI_COMM_NULL       ((MPI_Comm)0)
#define MPI_SUCCESS         0
#define MPI_UNDEFINED       (-32766)
#define MPI_ANY_SOURCE      (-1)
#define MPI_ANY_TAG         (-1)
#define MPI_STATUS_IGNORE   ((MPI_Status *)0)
#define MPI_THREAD_FUNNELED 1

#define MPI_INT8_T   1
#define MPI_UINT16_T 2
#define MPI_INT      4
#define MPI_INT32_T  4
#define MPI_FLOAT    4

int MPI_Init_thread(int * argc, char *** argv, int required, int * provided);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int * rank);
int MPI_Comm_size(MPI_Comm comm, int * size);
int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm * newcomm);
int MPI_Comm_free(MPI_Comm * comm);
int MPI_Buffer_attach(void * buf, int size);
int MPI_Buffer_detach(void * buf, int * size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Probe(int src, int tag, MPI_Comm comm, MPI_Status * st);
int MPI_Iprobe(int src, int tag, MPI_Comm comm, int * flag, MPI_Status * st);
int MPI_Get_count(const MPI_Status * st, MPI_Datatype dt, int * count);
int MPI_Recv(void * buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Status * st);
int MPI_Irecv(void * buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Request * req);
int MPI_Wait(MPI_Request * req, MPI_Status * st);
int MPI_Bsend(const void * buf, int count, MPI_Datatype dt, int dst, int tag, MPI_Comm comm);
int MPI_Bcast(void * buf, int count, MPI_Datatype dt, int root, MPI_Comm comm);
int MPI_Scatter(const void * sbuf, int scount, MPI_Datatype sdt, void * rbuf, int rcount, MPI_Datatype rdt,
                int root, MPI_Comm comm);

#ifdef __cplusplus
}
#endif
#endif
