"""One rank of the multi-GPU pipeline test (launched by tests/test_gpu_pipeline.py with torch.distributed.run, one rank
per GPU).  The layer-split target must generate exactly the tokens of the same model on one GPU, and every stage's
device replica of the KV bookkeeping must equal the head's."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from pipeinfer_b200 import llama, pipeline      # noqa: E402

Q4_K = 12
TGT = (8, 1024, 8, 2, 2048, 4096, Q4_K)          # layers, embd, heads, kv heads, ff, vocab, type
DFT = (2, 512, 8, 2, 1024, 4096, Q4_K)


def main():
    rank, world, dev = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(dev)
    dist.init_process_group("cpu:gloo,cuda:nccl")
    pipeline.open_pipeline(rank, world, dev)
    ranges = pipeline.split_layers(TGT[0], pipeline.default_weights(world, float(os.environ.get("PIB_TEST_HEAD_SHARE", "0.25"))))
    assert ranges[0][0] == 0 and ranges[-1][1] == TGT[0] and all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1)), ranges
    cap = 1000 * 2 + 2
    tgt = pipeline.StageContext(rank, world, ranges, 512, cap, synthetic=TGT + (1, 1, 1.0))
    prompt = [1] + [3 + (i * 7919) % (TGT[5] - 3) for i in range(40)]
    # full logits rows through the llama.h calls the reference's sampling code uses (llama_decode + llama_get_logits_ith):
    # on the head of a pipeline every row is fetched from the last stage
    full_logits = None
    if rank == 0:
        rc, full_logits = tgt.decode(prompt[:5], list(range(5)), [[0]] * 5)
        assert rc == 0 and full_logits is not None and np.isfinite(full_logits).all()
        tgt.L.llama_kv_cache_clear(tgt.ctx)
        tgt.pause()
    else:
        assert tgt.serve() == 0
    results = {}
    for n_draft, n_par in ((3, 2), (4, 1)):
        if rank == 0:
            dft = llama.Context(None, n_ctx=512, max_seq=cap, synthetic=DFT + (2, 1, 0.8))
            params = llama.spec_params(n_draft=n_draft, n_parallel=n_par, n_predict=48)
            toks, res, itl = llama.speculative_generate(tgt, dft, prompt, params)
            tgt.pause()
            results[(n_draft, n_par)] = (toks, res)
            dft.close()
        else:
            assert tgt.serve() == 0
        # every stage's device replica of the KV metadata == the head's
        st = tgt.kv_state(device=True)
        flat = torch.from_numpy(np.concatenate([[st["head"], st["used"], st["n"]], st["pos"], st["bits"].view(np.int32).ravel()]).astype(np.int64))
        alls = [torch.zeros_like(flat) for _ in range(world)]
        dist.all_gather(alls, flat)
        for r in range(world):
            assert torch.equal(alls[r], alls[0]), "device KV replica of stage %d differs from the head's" % r
        assert int(alls[0][1]) > 0
        if rank == 0:
            tgt.L.llama_kv_cache_clear(tgt.ctx)
    if rank == 0:
        tgt.shutdown()
        # the same model on this GPU alone: plain greedy decoding is the ground truth of speculative decoding
        one = llama.Context(None, n_ctx=512, max_seq=8, synthetic=TGT + (1, 1, 1.0))
        rc, lg1 = one.decode(prompt[:5], list(range(5)), [[0]] * 5)
        assert rc == 0 and np.array_equal(lg1, full_logits), float(np.abs(lg1 - full_logits).max())     # same kernels, same order
        one.L.llama_kv_cache_clear(one.ctx)
        ref, _, _ = llama.plain_generate(one, prompt, 48)
        for key, (toks, res) in results.items():
            n = min(len(ref), len(toks))
            assert n >= 48 and toks[:n] == ref[:n], (key, toks[:n], ref[:n])
            assert res["n_drafted"] > 0
        hop_bytes, hop_msgs, back = pipeline.traffic()
        assert hop_msgs > 0 and hop_bytes % (TGT[1] * 4) == 0 and back == 0
    else:
        assert tgt.serve() == 1
        hop_bytes, hop_msgs, back = pipeline.traffic()
        if rank == world - 1:
            # 80-byte acceptance records (verification batches: the walk runs on this stage), 96-byte top-8 rows (batches without a
            # table) and the five fetched logits rows
            assert hop_msgs == 0 and back > 5 * TGT[5] * 4 and (back - 5 * TGT[5] * 4) % 16 == 0, (hop_msgs, back)
    pipeline.close_pipeline()
    dist.destroy_process_group()
    print("PIPELINE_GPU_OK rank %d" % rank)


if __name__ == "__main__":
    main()
