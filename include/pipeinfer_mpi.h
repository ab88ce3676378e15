/*
 * pipeinfer_mpi.h -- C ABI of libpipeinfer_b200.so, transport level: the reference's ggml-mpi.h entry points
 * (ggml-mpi.h:70-301) mapped onto a shared-memory control ring (host records, in order, head -> every stage) and NCCL
 * send/recv over NVLink (device data), plus the calls that open that transport.  One process per GPU of one node.
 * Citations are to the reference repository.
 */
#ifndef PIPEINFER_MPI_H
#define PIPEINFER_MPI_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- opening the transport (replaces mpirun + MPI_Init_thread, ggml-mpi.c:58-82) ----
 * Every rank: pib_pipe_ring_open(name, rank, n, bytes) (rank 0 creates /dev/shm/<name>), then -- GPU runs only --
 * rank 0 draws two ids with pib_pipe_unique_id, the launcher's store (torch.distributed) hands them to all ranks, and
 * every rank calls pib_pipe_nccl_init collectively.  pib_pipe_barrier is a host barrier over the ring's ranks. */
int  pib_pipe_ring_open(const char * name, int rank, int n_ranks, size_t cap_bytes);
void pib_pipe_ring_close(void);
int  pib_pipe_barrier(void);
int  pib_pipe_rank(void);
int  pib_pipe_size(void);
int  pib_pipe_unique_id(uint8_t * out128);
int  pib_pipe_nccl_init(int rank, int size, const uint8_t * id_fwd, const uint8_t * id_back);
void pib_pipe_nccl_free(void);
/* all of the above from the launcher's environment (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_PORT): ring named after the port,
 * NCCL ids handed over through the ring's header page; returns the world size (1 = not a multi-process launch), < 0 on error */
int  pib_pipe_open_env(void);
/* bytes / messages of activations this rank sent to its next stage, and bytes of sampled rows it sent to the head */
void pib_pipe_traffic(uint64_t * hop_bytes, uint64_t * hop_msgs, uint64_t * back_bytes);
/* layer partition rule of ggml_mpi_split_range (ggml-mpi.c:523-559) without a context: out_pairs[2*i], [2*i+1] */
void pib_split_range(int n_stages, int start, int end, const float * weights, uint16_t * out_pairs);

/* ---- ggml-mpi.h ---- */
struct ggml_mpi_context;
void   ggml_mpi_backend_init(void);                                                        /* ggml-mpi.c:58  */
void   ggml_mpi_backend_free(void);                                                        /* :63  */
struct ggml_mpi_context * ggml_mpi_init(void);                                             /* :67  */
void   ggml_mpi_free(struct ggml_mpi_context * ctx);                                       /* :100 */
struct ggml_mpi_context * ggml_mpi_split_comm(struct ggml_mpi_context * ctx, int color, int key);   /* :84 */
int    ggml_mpi_rank(struct ggml_mpi_context * ctx);                                       /* :128 */
size_t ggml_mpi_size(struct ggml_mpi_context * ctx);                                       /* :132 */
void   ggml_mpi_barrier(struct ggml_mpi_context * ctx);                                    /* :136 */
int    ggml_mpi_next_node(struct ggml_mpi_context * ctx_mpi);                              /* :164 */
int    ggml_mpi_prev_node(struct ggml_mpi_context * ctx_mpi);                              /* :168 */
int    ggml_mpi_trans_id(struct ggml_mpi_context * ctx_mpi);                               /* :38  */
int    ggml_mpi_recv_trans_id(struct ggml_mpi_context * ctx_mpi);                          /* :42  */
void   ggml_mpi_inc_trans_id(struct ggml_mpi_context * ctx_mpi);                           /* :46  */
void   ggml_mpi_sync_ints_pipelined(struct ggml_mpi_context * ctx_mpi, int32_t * vals, int count, int tag);   /* :349 */
void   ggml_mpi_probe(struct ggml_mpi_context * ctx_mpi, int src, int tag);                /* :140 */
int    ggml_mpi_iprobe(struct ggml_mpi_context * ctx_mpi, int src, int tag);               /* :146 */
int    ggml_mpi_status_tag(struct ggml_mpi_context * ctx_mpi);                             /* :154 */
int    ggml_mpi_status_count_int32(struct ggml_mpi_context * ctx_mpi);                     /* :158 */
/* `val` is a DEVICE pointer; NCCL send / recv between any two ranks */
void   ggml_mpi_send_float_array_async(struct ggml_mpi_context * ctx_mpi, float * val, int arr_size, int dest, int tag);   /* :411 */
void   ggml_mpi_recv_float_array(struct ggml_mpi_context * ctx_mpi, float * val, int arr_size, int src, int tag);          /* :399 */
uint16_t ** ggml_mpi_split_range(struct ggml_mpi_context * ctx_mpi, uint16_t start, uint16_t end, float node_weights[]);   /* :523 */
void   ggml_mpi_scatter_layers(struct ggml_mpi_context * ctx_mpi, uint16_t ** layer_ranges);                               /* :561 */
/* additive: the range ggml_mpi_scatter_layers delivered to this rank */
void   ggml_mpi_layer_range(struct ggml_mpi_context * ctx_mpi, uint16_t * start, uint16_t * end);
struct ggml_cgraph;
void   ggml_mpi_synch_int(struct ggml_mpi_context * ctx_mpi, int32_t * val, int root);                                     /* :375, root 0 only */
void   ggml_mpi_synch_float(struct ggml_mpi_context * ctx_mpi, float * val, int root);                                     /* :387 */
void   ggml_mpi_sync_ints_pipelined_back(struct ggml_mpi_context * ctx_mpi, int32_t * vals, int count, int tag);           /* :361, no-op */
bool   ggml_mpi_is_decoding(struct ggml_mpi_context * ctx_mpi);                                                            /* :50 */
void   ggml_mpi_wait_recv(struct ggml_mpi_context * ctx_mpi);                                                              /* :516 */
/* the batch of a decode from the head to every stage, in the reference's message order (ggml-mpi.c:236-347) */
bool   ggml_mpi_eval_init(struct ggml_mpi_context * ctx_mpi, int32_t * n_tokens, int32_t ** tokens, int32_t ** pos, int32_t ** n_seq_ids,
                          int32_t *** seq_id, int8_t ** logits, int32_t * batch_id, bool receive_only);
/* graph hooks (:591, :683, :710): no ggml graph exists here, the hop is done inside the engine; kept so that callers link */
void   ggml_mpi_graph_creation_post(struct ggml_mpi_context * ctx_mpi, struct ggml_cgraph * cgraph, int n_layers);
bool   ggml_mpi_graph_compute_pre(struct ggml_mpi_context * ctx_mpi, struct ggml_cgraph * gf);
void   ggml_mpi_graph_compute_post(struct ggml_mpi_context * ctx_mpi, struct ggml_cgraph * gf);

/* ---- CPU-only test hooks: KV bookkeeping replicated through the control ring (no CUDA call) ---- */
void pib_kvtest_ship(void * kvtest, int enable);                 /* head: pib_kvtest_decode also publishes its record */
int  pib_kvtest_replay(void * kvtest);                            /* stage: apply records until the head pauses (0) */
int  pib_kvtest_pause(void * kvtest);                             /* head: flush the op-log, then pause the stages */

#ifdef __cplusplus
}
#endif
#endif
