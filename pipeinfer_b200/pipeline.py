"""Multi-GPU pipeline (one process per GPU, launched by torchrun): host-side mirror of what the reference does with
mpirun + --mpi-layer-split (llama.cpp:8901-8938, ggml-mpi.c).  torch.distributed is the launcher's plumbing only
(rendezvous, handing the shared-ring name and the NCCL ids around, barriers); the data path is libpipeinfer_b200.so.

    rank 0   head: scheduler + draft model + first layers of the target
    rank r   stage: layers [begin_r, end_r) of the target; the last one also owns the final norm, lm_head and sampling
"""
import ctypes as C
import os

import numpy as np

from . import llama
from .lib import PibError

RING_BYTES = 64 << 20


def _bind(L):
    if getattr(L, "_pib_pipe_bound", False):
        return
    L.pib_pipe_ring_open.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_size_t]
    L.pib_pipe_unique_id.argtypes = [C.c_void_p]
    L.pib_pipe_nccl_init.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.pib_pipe_traffic.argtypes = [C.POINTER(C.c_uint64)] * 3
    L.pib_split_range.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_uint16)]
    L.llama_pib_synthetic_model_slice.restype = C.c_void_p
    L.llama_pib_synthetic_model_slice.argtypes = [C.c_int] * 7 + [C.c_uint64, C.c_int, C.c_float] + [C.c_int] * 4
    L.llama_pib_load_model_slice.restype = C.c_void_p
    L.llama_pib_load_model_slice.argtypes = [C.c_char_p] + [C.c_int] * 4
    L.llama_pib_new_context_stage.restype = C.c_void_p
    L.llama_pib_new_context_stage.argtypes = [C.c_void_p, C.c_uint32, C.c_int32, C.c_int, C.c_int]
    L.llama_pib_stage_serve.argtypes = [C.c_void_p]
    L.llama_pib_pipe_pause.argtypes = [C.c_void_p]
    L.llama_pib_pipe_shutdown.argtypes = [C.c_void_p]
    L._pib_pipe_bound = True


def split_layers(n_layer, weights):
    """[(begin, end)] per stage by the rule of ggml_mpi_split_range (ggml-mpi.c:523-559) as llama_split_layers_weighted
    calls it (start 0, end n_layer - 1); the last stage runs to the end of the model (ggml-mpi.c:640-650)."""
    L = llama.api()
    _bind(L)
    n = len(weights)
    w = (C.c_float * n)(*weights)
    out = (C.c_uint16 * (2 * n))()
    L.pib_split_range(n, 0, n_layer - 1, w, out)
    r = [(int(out[2 * i]), int(out[2 * i + 1])) for i in range(n)]
    r[-1] = (r[-1][0], n_layer)
    return r


def default_weights(world, head_share=None):
    """the head also runs the draft model, so by default it takes a smaller share of the target's layers (the authors'
    launch script does the same: --mpi-layer-split 0.1,0.15,... scripts/dolphin_tinyllama_8_nodes.sh)"""
    if world == 1:
        return [1.0]
    if head_share is None:
        head_share = 0.75 / world
    rest = (1.0 - head_share) / (world - 1)
    return [head_share] + [rest] * (world - 1)


def open_pipeline(rank, world, device):
    """collective: control ring + the two NCCL communicators.  torch.distributed must be initialised (any backend that
    can move python objects; the gloo side is used)."""
    import torch
    import torch.distributed as dist
    L = llama.api()
    _bind(L)
    L.pib_init.argtypes = [C.c_int]
    if L.pib_init(device) != 0:
        raise PibError(L.pib_last_error().decode())
    cpu = torch.device("cpu")
    obj = [None]
    if rank == 0:
        ids = (C.c_uint8 * 256)()
        if L.pib_pipe_unique_id(ids) != 0 or L.pib_pipe_unique_id(C.byref(ids, 128)) != 0:
            raise PibError(L.pib_last_error().decode())
        obj = [("/pib_ring_%d" % os.getpid(), bytes(ids))]
    dist.broadcast_object_list(obj, src=0, device=cpu)
    name, ids = obj[0]
    if L.pib_pipe_ring_open(name.encode(), rank, world, RING_BYTES) != 0:
        raise PibError(L.pib_last_error().decode())
    buf = (C.c_uint8 * 256).from_buffer_copy(ids)
    if L.pib_pipe_nccl_init(rank, world, buf, C.byref(buf, 128)) != 0:
        raise PibError(L.pib_last_error().decode())
    return L


def close_pipeline():
    L = llama.api()
    L.pib_pipe_nccl_free()
    L.pib_pipe_ring_close()


class StageContext(llama.Context):
    """this rank's share of a target model + a context bound to its pipeline position"""

    def __init__(self, rank, world, ranges, n_ctx, max_seq, synthetic=None, model_path=None):
        self.L = L = llama.api()
        _bind(L)
        L.llama_backend_init(False)
        b, e = ranges[rank]
        first, last = int(rank == 0), int(rank == world - 1)
        if synthetic is not None:
            self.model = L.llama_pib_synthetic_model_slice(*synthetic, b, e, first, last)
        else:
            self.model = L.llama_pib_load_model_slice(model_path.encode(), b, e, first, last)
        if not self.model:
            raise PibError(L.pib_last_error().decode())
        self.ctx = L.llama_pib_new_context_stage(self.model, n_ctx, max_seq, rank, world)
        if not self.ctx:
            raise PibError(L.pib_last_error().decode())
        self.n_vocab = L.llama_n_vocab(self.model)
        self.n_ctx, self.max_seq, self._bid = n_ctx, max_seq, 0
        self.rank, self.world, self.layers = rank, world, (b, e)

    def serve(self):
        """ranks > 0: run the stage loop until the head pauses (0) or shuts the pipeline down (1)"""
        rc = self.L.llama_pib_stage_serve(self.ctx)
        if rc < 0:
            raise PibError(self.L.pib_last_error().decode())
        return rc

    def pause(self):
        if self.L.llama_pib_pipe_pause(self.ctx) != 0:
            raise PibError(self.L.pib_last_error().decode())

    def shutdown(self):
        self.L.llama_pib_pipe_shutdown(self.ctx)


def cancel_stats():
    """this stage's cancel-path latencies: (records, mean / max us head decision -> record seen by the stage's host, mean / max us
    head decision -> cancel word written on the stage's GPU; the latter only when PIB_CANCEL_TIMING=1)"""
    L = llama.api()
    n = C.c_uint64()
    v = [C.c_double() for _ in range(4)]
    L.pib_pipe_cancel_stats(C.byref(n), *[C.byref(x) for x in v])
    return (n.value,) + tuple(x.value for x in v)


def traffic():
    L = llama.api()
    _bind(L)
    a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
    L.pib_pipe_traffic(C.byref(a), C.byref(b), C.byref(c))
    return a.value, b.value, c.value
