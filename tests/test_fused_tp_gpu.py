"""Fused GEMM + all-reduce over NVLink peer memory (needs 2 GPUs): the kernel against an fp32 oracle, and a TP=2
transformer trained with the fused collectives against the same run on NCCL all-reduces."""
import os

import pytest
import torch

from tests.transformer_utils import tiny_config, write_dataset
from tests.utils import dist_launcher

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")]


def _kernel_worker(rank: int, world_size: int, master_port: int, two_shot: bool = False):
    if two_shot:  # read once by the library: force the owner-reduces / final-tile-broadcast protocol also at tp = 2
        import os

        os.environ["SCALING_B200_AR_TWO_SHOT"] = "1"
    import torch.distributed as dist

    from scaling_b200.ops import kernels as K
    from scaling_b200.parallel.symm import TPComm

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{master_port}", rank=rank, world_size=world_size,
                            device_id=torch.device("cuda", rank))
    comm = TPComm(dist.group.WORLD, torch.device("cuda", rank))
    worst = 0.0
    for (M, N, Kd, b_mn, has_bias) in [(512, 256, 128, False, False), (1000, 520, 264, False, True),
                                       (256, 128, 64, True, False), (2048, 1024, 1536, True, False),
                                       (4096, 4096, 512, False, False), (300, 136, 72, True, False)]:
        g = torch.Generator(device="cuda").manual_seed(7 + rank)
        a = torch.randn(M, Kd, device="cuda", generator=g).bfloat16()
        b = (torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda", generator=g) * 0.1).bfloat16()
        bias = torch.randn(N, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3)).bfloat16() if has_bias else None
        ref = a.float() @ (b.float() if b_mn else b.float().t())
        dist.all_reduce(ref)
        if bias is not None:
            ref = ref + bias.float()
        for _ in range(3):  # repeated calls exercise the parity double buffer and the epoch flags
            out = K.gemm_allreduce(a, b, comm, b_mn=b_mn, bias=bias)
        torch.cuda.synchronize()
        err = ((out.float() - ref).abs().max() / ref.abs().max()).item()
        gathered = [torch.empty_like(out) for _ in range(world_size)]
        dist.all_gather(gathered, out)
        assert all(torch.equal(gathered[0], x) for x in gathered), "ranks disagree bitwise"
        assert err < 2e-2, (M, N, Kd, err)
        worst = max(worst, err)
    return worst


def _rs_worker(rank: int, world_size: int, master_port: int):
    import torch.distributed as dist

    from scaling_b200.ops import kernels as K
    from scaling_b200.parallel.symm import TPComm

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{master_port}", rank=rank, world_size=world_size,
                            device_id=torch.device("cuda", rank))
    comm = TPComm(dist.group.WORLD, torch.device("cuda", rank))
    worst = 0.0
    for (batch, seq, N, Kd, b_mn, has_bias) in [(1, 512 * world_size, 256, 128, False, False),
                                                (2, 256 * world_size, 520, 192, False, True),
                                                (3, 256 * world_size, 384, 320, True, False)]:
        M = batch * seq
        g = torch.Generator(device="cuda").manual_seed(17 + rank)
        a = torch.randn(M, Kd, device="cuda", generator=g).bfloat16()
        b = (torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda", generator=g) * 0.1).bfloat16()
        bias = torch.randn(N, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3)).bfloat16() if has_bias else None
        full = a.float() @ (b.float() if b_mn else b.float().t())
        dist.all_reduce(full)
        if bias is not None:
            full = full + bias.float()
        per = seq // world_size
        want = full.view(batch, seq, N)[:, rank * per:(rank + 1) * per].reshape(batch * per, N)
        for _ in range(3):
            out = K.gemm_reduce_scatter(a, b, comm, seq, b_mn=b_mn, bias=bias)
        # interleave with the all-reduce flavour on the same workspace: epochs / parity are shared
        ar = K.gemm_allreduce(a, b, comm, b_mn=b_mn, bias=bias)
        torch.cuda.synchronize()
        assert out.shape == want.shape
        err = ((out.float() - want).abs().max() / want.abs().max()).item()
        err_ar = ((ar.float() - full).abs().max() / full.abs().max()).item()
        assert err < 2e-2 and err_ar < 2e-2, (batch, seq, N, Kd, err, err_ar)
        worst = max(worst, err)
    return worst


@pytest.mark.parametrize("world", [2, 4])
def test_gemm_reduce_scatter_kernel(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = dist_launcher(_rs_worker, world_size=world, timeout=300)
    assert len(res) == world and all(v < 2e-2 for v in res.values())


def _stress_worker(rank: int, world_size: int, master_port: int, iters: int):
    """Signal/wait protocols under skew (fused all-reduce, reduce-scatter tile owners, arrival-gated all-gather -> GEMM):
    ranks are delayed by random amounts before (and between) launches, shapes, layouts and collectives change from call to
    call, and results are checked against NCCL.  A lost flag, a stale-parity
    read or a premature buffer reuse shows up as a mismatch or as the kernel's 30 s watchdog trap."""
    import random

    import torch.distributed as dist

    from scaling_b200.ops import kernels as K
    from scaling_b200.parallel.symm import TPComm

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{master_port}", rank=rank, world_size=world_size,
                            device_id=torch.device("cuda", rank))
    comm = TPComm(dist.group.WORLD, torch.device("cuda", rank))
    shapes = [(256, 128, 64, False), (1024, 512, 256, True), (4096, 4096, 128, False), (768, 1280, 192, True),
              (2048, 256, 1024, False)]
    gen = torch.Generator(device="cuda").manual_seed(100 + rank)
    data = {}
    for (M, N, Kd, b_mn) in shapes:
        a = torch.randn(M, Kd, device="cuda", generator=gen).bfloat16()
        b = (torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda", generator=gen) * 0.1).bfloat16()
        want = K.gemm(a, b, b_mn=b_mn, out_dtype=torch.float32)
        dist.all_reduce(want)
        data[(M, N, Kd, b_mn)] = (a, b, want)
    order = random.Random(7)       # same call sequence on every rank
    jitter = random.Random(rank)   # different delays on every rank
    worst = 0.0
    for i in range(iters):
        shape = order.choice(shapes)
        a, b, want = data[shape]
        if jitter.random() < 0.5:
            torch.cuda._sleep(int(jitter.random() * 3_000_000))  # up to ~2 ms of device-side skew
        M = shape[0]
        seq = M // 2 if (M // 2) % (256 * world_size) == 0 else (M if M % (256 * world_size) == 0 else 0)
        mode = order.random()  # same choice on every rank: all-reduce, reduce-scatter or gated all-gather -> GEMM
        if mode < 0.3 and seq:
            # reduce-scatter flavour of the same kernel (tile owners): this rank's sequence slice of the sum
            out = K.gemm_reduce_scatter(a, b, comm, seq, b_mn=shape[3])
            sl = seq // world_size
            ref = want.view(M // seq, seq, -1)[:, rank * sl:(rank + 1) * sl].reshape(out.shape)
        elif mode < 0.5 and M % (256 * world_size) == 0 and not shape[3] and shape[1] >= 128:
            # all-gather -> GEMM: every rank contributes a row shard of A, the gated producer consumes it as it lands
            shard = a.view(1, M, -1)[:, rank * (M // world_size):(rank + 1) * (M // world_size)].contiguous()
            gathered = K.seq_gather_gated(shard, comm)
            out = K.gemm(gathered.view(M, -1), b, out_dtype=torch.float32)
            parts = [torch.empty_like(shard) for _ in range(world_size)]
            dist.all_gather(parts, shard)
            ref = torch.cat(parts, dim=1).view(M, -1).float() @ b.float().t()
        else:
            out = K.gemm_allreduce(a, b, comm, b_mn=shape[3])
            ref = want
        if i % 8 == 0 or i > iters - 4 or mode < 0.5:
            err = ((out.float() - ref).abs().max() / ref.abs().max()).item()
            assert err < 2e-2, (i, shape, mode, err)
            worst = max(worst, err)
    torch.cuda.synchronize()
    return worst


def test_gemm_allreduce_protocol_stress():
    res = dist_launcher(_stress_worker, world_size=2, timeout=400, iters=300)
    assert all(v < 2e-2 for v in res.values())


@pytest.mark.parametrize("world", [2, 4])
def test_gemm_allreduce_kernel(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = dist_launcher(_kernel_worker, world_size=world, timeout=300)
    assert len(res) == world and all(v < 2e-2 for v in res.values())


def test_gemm_allreduce_two_shot_protocol_on_two_ranks():
    """The tp > 2 protocol (partials to the tile's owner, owner stores the final tile into the peers' slots, peers copy it
    out) forced on two ranks, so that it is covered wherever two GPUs are."""
    res = dist_launcher(_kernel_worker, world_size=2, timeout=300, two_shot=True)
    assert len(res) == 2 and all(v < 2e-2 for v in res.values())


def _train_worker(rank: int, world_size: int, master_port: int, config_dict, fused: bool):
    from scaling_b200.core import LaunchConfig
    from scaling_b200.ops import _lib
    from scaling_b200.parallel import symm
    from scaling_b200.transformer.train import main

    symm.set_fused_tp(fused)
    n0 = _lib.launch_count()
    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=config_dict)
    metrics = main(launch, return_metrics=True)
    assert _lib.launch_count() > n0
    fused_calls = sum(c.calls for c in symm._COMMS.values())
    assert (fused_calls > 0) == fused, f"fused collective kernels ran {fused_calls} times with fused={fused}"
    if config_dict["topology"].get("sequence_parallel"):
        from scaling_b200.ops import kernels as K

        # sequence parallel + peer memory: the all-gathers ran INSIDE the q|k|v and gate|up GEMMs (gated producer)
        assert bool(K._gates) == fused, f"gated all-gather -> GEMM buffers: {len(K._gates)} with fused={fused}"
    return [m["training/loss"] for m in (metrics or [])]


@pytest.mark.parametrize("arch", [
    {"num_attention_heads": 4, "attention_num_kv_heads": 2},
    # GPT-style: biases (added once, after the fused reduction), fused qkv, GELU MLP, LayerNorm, tied LM head
    {"num_attention_heads": 4, "attention_num_kv_heads": None, "attention_qkv_in_one": True, "norm_type": "layernorm",
     "mlp_type": "default", "mlp_factor": 4.0, "attention_bias": True, "mlp_bias": True, "weight_tying": True,
     "relative_position_embedding_type": "rotary"},
])
def test_tp2_training_fused_matches_nccl(tmp_path, arch):
    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=512, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 512, "sequence_length": 256, "precision": "bfloat16",
         "masked_softmax": {"kernel": "flash_attention"}}
    a.update(arch)
    cfg = tiny_config(prefix, mp=2, iters=5, gas=2, arch=a)
    fused = dist_launcher(_train_worker, world_size=2, timeout=400, config_dict=cfg, fused=True)[0]
    plain = dist_launcher(_train_worker, world_size=2, timeout=400, config_dict=cfg, fused=False)[0]
    assert len(fused) == 5 and all(x == x for x in fused)
    for x, y in zip(fused, plain):
        assert abs(x - y) / abs(y) < 2e-2, (fused, plain)


def _pp_tp_worker(rank: int, world_size: int, master_port: int, config_dict):
    from scaling_b200.core import LaunchConfig
    from scaling_b200.transformer.train import main

    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=config_dict)
    metrics = main(launch, return_metrics=True)
    return [m["training/loss"] for m in (metrics or [])]


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs 4 GPUs")
def test_tp2_pp2_training_on_nccl(tmp_path):
    """The headline layout in small: tensor x pipeline parallel with the fused collectives, evaluation in between."""
    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=512, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 512, "num_attention_heads": 4, "attention_num_kv_heads": 2,
         "sequence_length": 256, "precision": "bfloat16", "num_layers": 4, "masked_softmax": {"kernel": "flash_attention"}}
    cfg = tiny_config(prefix, mp=2, pp=2, iters=6, gas=4, arch=a)
    losses = dist_launcher(_pp_tp_worker, world_size=4, timeout=400, config_dict=cfg)[0]
    assert len(losses) == 6 and all(x == x for x in losses), losses
    # same model on plain TP=2 (no pipeline): the two layouts share the initial weights (seeded per (pp, dp) cell
    # differently, so only the scale of the curve is compared) — the pipeline run must not diverge
    assert max(losses) < 1.5 * losses[0] and min(losses) > 0.3 * losses[0], losses


def _zero_worker(rank: int, world_size: int, master_port: int, config_dict, fused: bool):
    from scaling_b200.core import LaunchConfig
    from scaling_b200.parallel import symm
    from scaling_b200.transformer.train import main

    symm.set_fused_zero(fused)
    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=config_dict)
    metrics = main(launch, return_metrics=True)
    return [(m["training/loss"], m["training/global_grad_norm"]) for m in (metrics or [])]


def test_dp2_fused_zero_matches_nccl(tmp_path):
    """ZeRO-1 with the reduce-scatter as peer loads (sb_zero_reduce) and the all-gather as peer stores
    (sb_adamw_zero) against the NCCL reduce_scatter / all_gather path: two summands commute, so the runs agree bitwise."""
    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=512, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 256, "num_attention_heads": 2, "attention_num_kv_heads": 1,
         "sequence_length": 128, "precision": "bfloat16", "masked_softmax": {"kernel": "flash_attention"}}
    cfg = tiny_config(prefix, dp=2, iters=6, gas=2, arch=a)
    fused = dist_launcher(_zero_worker, world_size=2, timeout=400, config_dict=cfg, fused=True)[0]
    plain = dist_launcher(_zero_worker, world_size=2, timeout=400, config_dict=cfg, fused=False)[0]
    assert len(fused) == 6
    assert fused == plain, (fused, plain)
    assert fused[-1][0] < fused[0][0]


def test_dp2_fused_allreduce_without_zero_matches_nccl(tmp_path):
    """Plain data parallelism (optimizer.zero = false): the gradient all-reduce as ONE kernel of peer loads + peer
    stores (sb_dp_allreduce) against NCCL all_reduce — two summands commute, so the runs agree bitwise."""
    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=512, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 256, "num_attention_heads": 2, "attention_num_kv_heads": 1,
         "sequence_length": 128, "precision": "bfloat16", "masked_softmax": {"kernel": "flash_attention"}}
    cfg = tiny_config(prefix, dp=2, iters=6, gas=2, arch=a, optimizer={"zero": False})
    fused = dist_launcher(_zero_worker, world_size=2, timeout=400, config_dict=cfg, fused=True)[0]
    plain = dist_launcher(_zero_worker, world_size=2, timeout=400, config_dict=cfg, fused=False)[0]
    assert len(fused) == 6
    assert fused == plain, (fused, plain)
    assert fused[-1][0] < fused[0][0]


def _dp_kernel_stress_worker(rank, world_size, master_port):
    """Device barrier + peer reduce / all-reduce kernels under random per-rank launch skew, many iterations."""
    import random

    import torch
    import torch.distributed as dist

    from scaling_b200.ops import _lib as native
    from scaling_b200.parallel.symm import SymmetricArena

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{master_port}", rank=rank, world_size=world_size,
                            device_id=torch.device("cuda", rank))
    dev = torch.device("cuda", rank)
    n = 3 * 1024 * 128 * world_size + 128 * world_size
    arena = SymmetricArena(n, torch.float32, dist.group.WORLD, dev)
    shard = n // world_size
    own = torch.empty(shard, device=dev)
    rng = random.Random(1234 + rank)
    spin = torch.empty(1 << 22, device=dev)
    ok = True
    for it in range(60):
        g = torch.Generator(device="cuda").manual_seed(1000 * it + rank)
        local = torch.randn(n, device=dev, generator=g)
        arena.tensor.copy_(local)
        for _ in range(rng.randrange(0, 4)):  # random skew: this rank reaches the barrier late
            spin.normal_()
        expect = torch.zeros(n, device=dev)
        for r in range(world_size):
            gr = torch.Generator(device="cuda").manual_seed(1000 * it + r)
            expect += torch.randn(n, device=dev, generator=gr)
        arena.barrier()
        if it % 2 == 0:  # ZeRO flavour: reduce the own shard out of every arena
            native.call("sb_zero_reduce", arena.ptr_array(rank * shard), world_size, native.ptr(own), shard, 1)
            arena.barrier()
            ok = ok and torch.equal(own, expect[rank * shard:(rank + 1) * shard])
        else:  # plain DP flavour: all-reduce in place
            native.call("sb_dp_allreduce", arena.ptr_array(rank * shard), world_size, shard, 1)
            arena.barrier()
            ok = ok and torch.equal(arena.tensor, expect)
        arena.barrier()  # nobody overwrites its arena while a peer still reads it
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()
    return bool(ok)


def test_dp_reduce_kernels_and_device_barrier_under_launch_skew():
    assert all(dist_launcher(_dp_kernel_stress_worker, world_size=2, timeout=300).values())


def _lora_worker(rank: int, world_size: int, master_port: int, config_dict, fused: bool):
    from scaling_b200.core import LaunchConfig
    from scaling_b200.ops import functional as OF
    from scaling_b200.transformer.train import main

    OF.set_lora_fusion(fused)
    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=config_dict)
    metrics = main(launch, return_metrics=True)
    return [m["training/loss"] for m in (metrics or [])]


@pytest.mark.parametrize("mp", [1, 2])
def test_lora_fused_projections_match_separate_lora_linears(tmp_path, mp):
    """LoRA on q/k/v/dense trained with the low-rank pair inside the projection GEMMs (extra K-slab, and for the output
    projection inside the fused all-reduce, no activation gather) against base GEMM + separate LoRA linears."""
    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=1024, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 512, "num_attention_heads": 4, "attention_num_kv_heads": 2,
         "sequence_length": 256, "precision": "bfloat16", "masked_softmax": {"kernel": "flash_attention"},
         "lora_config": {"name": "lo", "rank": 16, "alpha": 32}}
    cfg = tiny_config(prefix, mp=mp, iters=6, gas=2, arch=a,
                      learning_rate_scheduler={"learning_rate": 0.002, "learning_rate_decay_style": "constant"})
    fused = dist_launcher(_lora_worker, world_size=mp, timeout=400, config_dict=cfg, fused=True)[0]
    plain = dist_launcher(_lora_worker, world_size=mp, timeout=400, config_dict=cfg, fused=False)[0]
    assert len(fused) == 6 and all(x == x for x in fused) and fused[-1] < fused[0], (fused, plain)
    for x, y in zip(fused, plain):
        assert abs(x - y) / abs(y) < 2e-2, (fused, plain)


@pytest.mark.parametrize("sp", [False, True])
def test_tp2_loss_and_gradient_norm_match_tp1_from_the_same_checkpoint(tmp_path, sp):
    """Gradient parity across layouts on the peer-memory kernels: a TP=1 run writes a checkpoint; TP=1 and TP=2 (fused
    all-reduce, or sequence parallel with gated gathers / fused reduce-scatters) resume from it on the same data.  The
    first step's loss and GLOBAL GRADIENT NORM (every gradient of the model, before any update) must agree."""
    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=1024, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 512, "num_attention_heads": 4, "attention_num_kv_heads": 2,
         "sequence_length": 512, "precision": "bfloat16", "masked_softmax": {"kernel": "flash_attention"}}
    ckpt = tmp_path / "ckpt"
    base = dict(iters=1, gas=2, arch=a, learning_rate_scheduler={"learning_rate": 1e-4, "learning_rate_decay_style": "constant"})
    dist_launcher(_zero_worker, world_size=1, timeout=400, fused=True,
                  config_dict=tiny_config(prefix, trainer={"save_dir": str(ckpt), "save_interval": 1}, **base))
    load = {"load_dir": str(ckpt), "load_optimizer_states": False, "load_context": False, "assert_checkpoint_loaded": True}
    base["iters"] = 2
    one = dist_launcher(_zero_worker, world_size=1, timeout=400, fused=True, config_dict=tiny_config(prefix, trainer=load, **base))[0]
    two = dist_launcher(_zero_worker, world_size=2, timeout=400, fused=True,
                        config_dict=tiny_config(prefix, mp=2, trainer=load, topology={"sequence_parallel": sp}, **base))[0]
    (l1, g1), (l2, g2) = one[0], two[0]
    assert abs(l1 - l2) / abs(l1) < 2e-3, (one, two)
    assert abs(g1 - g2) / abs(g1) < 2e-2, (one, two)


def test_sequence_parallel_fused_reduce_scatter_matches_nccl(tmp_path):
    """Sequence-parallel TP=2 training with the row-parallel reduce-scatters inside the GEMM launch against the same
    run on NCCL reduce-scatters."""
    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=1024, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 512, "num_attention_heads": 4, "attention_num_kv_heads": 2,
         "sequence_length": 512, "precision": "bfloat16", "masked_softmax": {"kernel": "flash_attention"}}
    cfg = tiny_config(prefix, mp=2, iters=5, gas=2, arch=a, topology={"sequence_parallel": True})
    fused = dist_launcher(_train_worker, world_size=2, timeout=400, config_dict=cfg, fused=True)[0]
    plain = dist_launcher(_train_worker, world_size=2, timeout=400, config_dict=cfg, fused=False)[0]
    assert len(fused) == 5 and all(x == x for x in fused)
    for x, y in zip(fused, plain):
        assert abs(x - y) / abs(y) < 2e-2, (fused, plain)
