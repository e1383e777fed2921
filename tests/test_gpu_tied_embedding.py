"""b200sft_embed_grad_rows (csrc/embed_grad.cu) and the tied-embedding gradient split on the GPU.

The kernel takes raw device pointers per rank, so `world` ranks are emulated on ONE GPU with `world` buffers
(as tests/test_gpu_allreduce_kernel_one_gpu.py does for the reduce kernel); the end-to-end case runs a world-size-1
NCCL DistributedDataParallel around a tiny tied SmolLM3 and compares the split path with stock DDP."""
import ctypes
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import sft_oracle as oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _call(lib, ids, dys, ns, cap, V, H, scale, grad, dev):
    world = len(ids)
    ws_bytes = lib.b200sft_embed_grad_workspace_bytes(world, cap, V, H)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    tab = lambda ts: (ctypes.c_uint64 * world)(*[t.data_ptr() for t in ts])
    rc = lib.b200sft_embed_grad_rows(tab(ids), tab(dys), tab(ns), world, cap, V, H, scale, grad.data_ptr(), ws.data_ptr(),
                                     ws_bytes, torch.cuda.current_stream(dev).cuda_stream)
    assert rc == 0, lib.b200sft_last_error().decode()
    torch.cuda.synchronize(dev)


def _case(world, cap, V, H, seed, dev, hot_rows=5, hot_frac=0.5):
    g = torch.Generator().manual_seed(seed)
    ids, dys, ns = [], [], []
    for r in range(world):
        n = cap if r % 2 == 0 else max(1, cap - 37 - r)              # ragged valid lengths
        i = torch.randint(0, V, (cap,), generator=g, dtype=torch.int32)
        hot = torch.rand(cap, generator=g) < hot_frac                # a few frequent tokens: long segments
        i[hot] = torch.randint(0, hot_rows, (int(hot.sum()),), generator=g, dtype=torch.int32)
        i[torch.rand(cap, generator=g) < 0.1] = -1                   # padding positions
        i[-1] = V + 5                                                # out of range: ignored
        ids.append(i.to(dev))
        dys.append((torch.randn(cap, H, generator=g) * 0.05).to(torch.bfloat16).to(dev))
        ns.append(torch.tensor(n, dtype=torch.int32, device=dev))
    grad = (torch.randn(V, H, generator=g) * 0.02).to(torch.bfloat16).to(dev)
    return ids, dys, ns, grad


@pytest.mark.parametrize("world,cap,V,H", [(1, 200, 97, 64), (2, 1024, 512, 256), (3, 1500, 1000, 128), (4, 2048, 4096, 2048),
                                           (8, 640, 300, 8192)])
def test_rows_kernel_bit_exact_against_the_restatement(cuda_device, world, cap, V, H):
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    ids, dys, ns, grad = _case(world, cap, V, H, seed=world * 7 + H, dev=cuda_device)
    before = grad.clone()
    scale = 1.0 / world
    _call(lib, ids, dys, ns, cap, V, H, scale, grad, cuda_device)
    ids_np = []
    for r in range(world):
        a = ids[r].cpu().numpy().copy()
        a[int(ns[r]):] = -1
        ids_np.append(a)
    want = oracle.embed_grad_rows_numpy(before.float().cpu().numpy(), ids_np, [d.float().cpu().numpy() for d in dys], scale)
    got = grad.float().cpu().numpy()
    assert np.array_equal(got, want)
    touched = np.unique(np.concatenate([a[(a >= 0) & (a < V)] for a in ids_np]))
    untouched = np.setdiff1d(np.arange(V), touched)
    assert np.array_equal(got[untouched], before.float().cpu().numpy()[untouched])      # other rows are not written


def test_rows_kernel_full_size_against_torch_embedding_backward(cuda_device):
    """The bench shape: 8 ranks x 8192 tokens, V = 128,256, H = 2048.  Against the literal reference ops on the same
    GPU (torch's embedding backward per rank, averaged) to bf16 round-off, and run twice: identical bits."""
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    world, cap, V, H = 8, 8192, 128256, 2048
    ids, dys, ns, grad = _case(world, cap, V, H, seed=11, dev=cuda_device, hot_rows=50, hot_frac=0.3)
    before = grad.clone()
    _call(lib, ids, dys, ns, cap, V, H, 1.0 / world, grad, cuda_device)
    again = before.clone()
    _call(lib, ids, dys, ns, cap, V, H, 1.0 / world, again, cuda_device)
    assert torch.equal(grad, again)
    exact = before.double()
    for r in range(world):
        n = int(ns[r])
        i = ids[r][:n].long()
        keep = (i >= 0) & (i < V)
        exact.index_add_(0, i[keep], dys[r][:n][keep].double() / world)
    scale = float(exact.abs().max())
    err = float((grad.double() - exact).abs().max()) / scale
    ref = before.clone()
    for r in range(world):                      # the reference's own arithmetic: bf16 embedding grads, bf16 average
        n = int(ns[r])
        i = ids[r][:n].long()
        i = torch.where((i >= 0) & (i < V), i, torch.full_like(i, V))      # park skipped tokens on a padding row
        ref += oracle.embedding_weight_grad(i, dys[r][:n], V + 1, padding_idx=V)[:V].div(world)
    err_ref = float((ref.double() - exact).abs().max()) / scale
    print(f"full size: max-rel-err vs exact {err:.3e} (reference bf16 path {err_ref:.3e})")
    assert err <= 2.0 ** -8 and err <= err_ref


def test_rows_kernel_rejects_bad_arguments(cuda_device):
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    t = torch.zeros(64, dtype=torch.int32, device=cuda_device)
    one = (ctypes.c_uint64 * 1)(t.data_ptr())
    s = torch.cuda.current_stream(cuda_device).cuda_stream
    assert lib.b200sft_embed_grad_rows(one, one, one, 1, 8, 16, 12, 1.0, t.data_ptr(), t.data_ptr(), 1 << 20, s) != 0   # H % 8
    assert lib.b200sft_embed_grad_rows(one, one, one, 17, 8, 16, 16, 1.0, t.data_ptr(), t.data_ptr(), 1 << 20, s) != 0  # world
    assert lib.b200sft_embed_grad_rows(one, one, one, 1, 8, 16, 16, 1.0, t.data_ptr(), t.data_ptr(), 16, s) == -3       # scratch
    assert b"workspace" in lib.b200sft_last_error()
    assert lib.b200sft_embed_grad_rows(None, one, one, 1, 8, 16, 16, 1.0, t.data_ptr(), t.data_ptr(), 1 << 20, s) != 0


def _ddp_worker(rank, port, ret, world=1):
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from torch.nn.parallel import DistributedDataParallel as DDP
    from transformers import SmolLM3Config, SmolLM3ForCausalLM

    from llm_fine_tune_distributed_b200.comm_hook import register_b200_comm_hook
    from llm_fine_tune_distributed_b200.loss import FusedCELossMixin
    from llm_fine_tune_distributed_b200.tied_embedding import TiedEmbeddingExchange
    from llm_fine_tune_distributed_b200.training import apply_freeze_policy

    def run(split):
        cfg = SmolLM3Config(vocab_size=4096, hidden_size=256, intermediate_size=512, num_hidden_layers=3,
                            num_attention_heads=4, num_key_value_heads=2, tie_word_embeddings=True, pad_token_id=9)
        cfg._attn_implementation = "sdpa"
        cfg.use_cache = False
        torch.manual_seed(0)
        model = SmolLM3ForCausalLM(cfg).to(torch.bfloat16).to(dev)
        apply_freeze_policy(model)
        ddp = DDP(model, device_ids=[rank], bucket_cap_mb=1)

        class Seam(FusedCELossMixin):
            num_processes = world
            model_accepts_loss_kwargs = True
            args = None
            b200sft_register_comm_hook = False

        seam = Seam()
        state = register_b200_comm_hook(ddp, nvlink_kernel=True)
        ex = TiedEmbeddingExchange(ddp, max_tokens=4 * 64) if split else None
        if ex is not None:
            ddp._b200sft_tied_exchange = ex
        g = torch.Generator().manual_seed(5 + rank)
        for it in range(3):
            a = torch.randint(0, 4096, (4, 64), generator=g).to(dev)
            b = torch.randint(0, 200, (4, 64), generator=g).to(dev)
            mask = torch.ones_like(b)
            mask[2, 50:] = 0
            n_items = (a.ne(-100).sum() + mask.sum()).to(dev)
            if world > 1:
                dist.all_reduce(n_items)
            with ddp.no_sync():
                seam.compute_loss(ddp, {"input_ids": a, "labels": a}, num_items_in_batch=n_items).backward()
            lab = torch.where(mask.bool(), b, torch.full_like(b, -100))
            seam.compute_loss(ddp, {"input_ids": b, "attention_mask": mask, "labels": lab}, num_items_in_batch=n_items).backward()
        torch.cuda.synchronize()
        return {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}, (ex.exchanges if ex else 0), state

    stock, _, _ = run(False)
    ours, exchanges, state = run(True)
    tied = "model.embed_tokens.weight"
    a, b = stock[tied].float(), ours[tied].float()
    gathered = [torch.zeros_like(b) for _ in range(world)]
    dist.all_gather(gathered, b.contiguous())
    ret[rank] = {"exchanges": exchanges, "others_equal": all(torch.equal(stock[k], ours[k]) for k in stock if k != tied),
              "replicas_identical": all(torch.equal(gathered[0], t) for t in gathered[1:]),
              "rel_err": float((a - b).abs().max() / a.abs().max()),
              "cos": float((a.flatten().double() @ b.flatten().double()) / (a.double().norm() * b.double().norm())),
              "path": state.path_used}
    dist.destroy_process_group()


def test_split_under_world1_nccl_ddp_matches_stock(cuda_device):
    ret = mp.Manager().dict()
    mp.spawn(_ddp_worker, args=(29691, ret), nprocs=1, join=True)
    r = ret[0]
    print(f"tied gradient, split path vs stock DDP (world 1, fused loss, NVLink hook {r['path']}): "
          f"max-rel-err {r['rel_err']:.3e}, cos {r['cos']:.6f}")
    assert r["exchanges"] == 3 and r["others_equal"]
    assert r["cos"] >= 0.9999 and r["rel_err"] <= 2.0 ** -6


def test_split_under_two_gpu_nccl_ddp_matches_stock():
    """Two real ranks: rows cross NVLink through the peers' symmetric-memory pointers."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ret = mp.Manager().dict()
    mp.spawn(_ddp_worker, args=(29693, ret, 2), nprocs=2, join=True)
    for rank in range(2):
        r = ret[rank]
        print(f"rank {rank}: tied gradient, split path vs stock DDP (2 GPUs, NVLink hook {r['path']}): "
              f"max-rel-err {r['rel_err']:.3e}, cos {r['cos']:.6f}")
        assert r["exchanges"] == 3 and r["others_equal"] and r["replicas_identical"]
        assert r["cos"] >= 0.9999 and r["rel_err"] <= 2.0 ** -6
