"""Per-rank bodies for the multi-GPU tests (one process per GPU, gloo bootstrap, SymmComm collectives)."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def _init(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)


def _gather_cpu(t, world):
    out = [torch.empty_like(t.cpu()) for _ in range(world)]
    dist.all_gather(out, t.cpu())
    return out


def _same(comm, got, expect, what):
    """Peer-load kernels sum in fixed rank order in fp32: bit-identical to the sequential oracle.  The NVLS kernels
    sum inside the switch (fp32, order chosen by the hardware): equal to the oracle up to one rounding step."""
    if comm.nvls_active:
        tol = 2 ** -7 if expect.dtype == torch.bfloat16 else 1e-6
        assert torch.allclose(got.float(), expect.float(), rtol=tol, atol=tol * 4), \
            (what, (got.float() - expect.float()).abs().max())
    else:
        assert torch.equal(got, expect), (what, (got.float() - expect.float()).abs().max())


def collectives(rank, world, port, out_dir, nvls="auto"):
    if nvls != "auto":
        os.environ["DTB_COMM_NVLS"] = str(nvls)         # read once by the library, before its first collective
    from distributed_training_b200.parallel.symm import SymmComm

    _init(rank, world, port)
    comm = SymmComm(torch.device("cuda", rank), segment_bytes=256 << 20)
    try:
        dev = comm.device
        comm.reserve(1 << 20)
        if nvls == "1" and not comm.multicast:
            torch.save({"ok": True, "skipped": "no multicast support on this box"},
                       os.path.join(out_dir, f"ok{rank}.pt"))
            return
        assert comm.nvls_active == (comm.multicast and (nvls == "1" or (nvls == "auto" and world >= 4)))
        # ---- all-reduce: one-shot and two-shot sizes, both dtypes, repeated (flag reuse) ----
        for dtype in (torch.float32, torch.bfloat16):
            for n in (21, 1000, 16 * 1024, 300_000, 4_000_000 + 8):
                buf = comm.alloc(n, dtype)
                for rep in range(3):
                    g = torch.Generator().manual_seed(1000 * rep + n)
                    per_rank = [torch.randn(n, generator=g).to(dtype) for _ in range(world)]
                    buf.copy_(per_rank[rank])
                    comm.all_reduce_avg_(buf).wait()
                    torch.cuda.synchronize()
                    acc = torch.zeros(n)
                    for r in range(world):                    # fixed rank order, fp32, like the kernel
                        acc += per_rank[r].float()
                    expect = (acc / world).to(dtype)
                    if n * buf.element_size() <= 64 << 10:        # one-shot (peer loads): always exact
                        assert torch.equal(buf.cpu(), expect), (dtype, n, rep)
                    else:
                        _same(comm, buf.cpu(), expect, (dtype, n, rep))
                    got = _gather_cpu(buf, world)
                    assert all(torch.equal(got[0], t) for t in got), "ranks disagree bitwise"
        # ---- copy-engine path (what large buckets take): same contract; threshold lowered so small buffers use it ----
        comm.copy_engine, comm.copy_engine_min_bytes = True, 1 << 16
        for n in (world * 16 * 4096, world * 16 * 100_000):
            comm.reserve(2 * n * 2 + 4096)
            buf, par = comm.alloc(n, torch.bfloat16), comm.alloc(n, torch.bfloat16)
            assert comm._ce_scratch(n * 2) is not None
            for rep in range(3):
                g = torch.Generator().manual_seed(77 * rep + n)
                per_rank = [torch.randn(n, generator=g).to(torch.bfloat16) for _ in range(world)]
                p0 = torch.randn(n, generator=g).to(torch.bfloat16)
                acc = torch.zeros(n)
                for r in range(world):
                    acc += per_rank[r].float()
                mean = acc / world
                buf.copy_(per_rank[rank])
                comm.all_reduce_avg_(buf).wait()
                torch.cuda.synchronize()
                assert torch.equal(buf.cpu(), mean.to(torch.bfloat16)), ("ce mean", n, rep)
                # fused SGD: gradients stay untouched, every rank ends with the same new parameters
                buf.copy_(per_rank[rank])
                par.copy_(p0)
                comm.all_reduce_sgd_(buf, par, 0.05, 0.01).wait()
                torch.cuda.synchronize()
                expect = p0.float() * (1 - 0.05 * 0.01) - 0.05 * mean
                assert torch.allclose(par.cpu().float(), expect, rtol=2 ** -7, atol=1e-6), ("ce sgd", n, rep)
                assert torch.equal(buf.cpu(), per_rank[rank]), "gradients must be left untouched"
                got = _gather_cpu(par, world)
                assert all(torch.equal(got[0], t) for t in got), "ranks disagree bitwise after ce sgd"
        comm.copy_engine_min_bytes = 16 << 20
        comm.copy_engine = False
        # ---- exclusive mode (whole-SM 1024-thread CTAs in clusters of two: what large DDP buckets take while backward
        # GEMMs leave SMs free): same bitwise contract as the co-resident kernels ----
        comm.exclusive_sms, comm.exclusive_now, comm.exclusive_min_bytes = 4, True, 1 << 16
        for n in (world * 16 * 4096, world * 16 * 100_000 + 16 * 8):
            comm.reserve(2 * n * 2 + 4096)
            buf, par = comm.alloc(n, torch.bfloat16), comm.alloc(n, torch.bfloat16)
            assert comm._ctas_for(n * 2) & (1 << 20)
            for rep in range(2):
                g = torch.Generator().manual_seed(91 * rep + n)
                per_rank = [torch.randn(n, generator=g).to(torch.bfloat16) for _ in range(world)]
                p0 = torch.randn(n, generator=g).to(torch.bfloat16)
                acc = torch.zeros(n)
                for r in range(world):
                    acc += per_rank[r].float()
                mean = acc / world
                buf.copy_(per_rank[rank])
                comm.all_reduce_avg_(buf).wait()
                torch.cuda.synchronize()
                _same(comm, buf.cpu(), mean.to(torch.bfloat16), ("exclusive mean", n, rep))
                buf.copy_(per_rank[rank])
                par.copy_(p0)
                comm.all_reduce_sgd_(buf, par, 0.05, 0.01).wait()
                torch.cuda.synchronize()
                expect = p0.float() * (1 - 0.05 * 0.01) - 0.05 * mean
                # NVLS hands back the bf16-rounded gradient sum (peer loads keep it in fp32 up to the update)
                atol = 1e-3 if comm.nvls_active else 1e-6
                assert torch.allclose(par.cpu().float(), expect, rtol=2 ** -7, atol=atol), ("exclusive sgd", n, rep)
                got = _gather_cpu(par, world)
                assert all(torch.equal(got[0], t) for t in got), "ranks disagree bitwise after exclusive sgd"
        comm.exclusive_now = False
        # ---- co-resident fused all-reduce + SGD (what small/medium DDP buckets take) ----
        for n in (16 * 4096, world * 16 * 50_000 + 64):
            comm.reserve(2 * n * 2 + 4096)
            buf, par = comm.alloc(n, torch.bfloat16), comm.alloc(n, torch.bfloat16)
            g = torch.Generator().manual_seed(17 + n)
            per_rank = [torch.randn(n, generator=g).to(torch.bfloat16) for _ in range(world)]
            p0 = torch.randn(n, generator=g).to(torch.bfloat16)
            mean = sum(t.float() for t in per_rank) / world
            buf.copy_(per_rank[rank])
            par.copy_(p0)
            comm.all_reduce_sgd_(buf, par, 0.05, 0.01).wait()
            torch.cuda.synchronize()
            expect = p0.float() * (1 - 0.05 * 0.01) - 0.05 * mean
            atol = 1e-3 if comm.nvls_active else 1e-6
            assert torch.allclose(par.cpu().float(), expect, rtol=2 ** -7, atol=atol), ("sgd", n)
            assert torch.equal(buf.cpu(), per_rank[rank]), "gradients must be left untouched"
            got = _gather_cpu(par, world)
            assert all(torch.equal(got[0], t) for t in got), "ranks disagree bitwise after fused sgd"
        # ---- broadcast ----
        t1 = torch.full((1000, 33), float(rank + 1), device=dev)
        t2 = torch.full((7,), float(rank + 5), device=dev, dtype=torch.bfloat16)
        comm.broadcast_([t1, t2], src=0)
        assert float(t1.min()) == 1.0 and float(t1.max()) == 1.0 and float(t2.float().max()) == 5.0
        # ---- all-gather / reduce-scatter ----
        n = 100_000
        comm.reserve((world + 2) * n * 4 + 4096)
        full = comm.alloc(world * n, torch.bfloat16)
        shard = comm.alloc(n, torch.bfloat16)
        shard.fill_(rank + 1)
        comm.all_gather_into(full, shard)
        torch.cuda.synchronize()
        for r in range(world):
            assert float(full[r * n:(r + 1) * n].float().min()) == r + 1 == float(full[r * n:(r + 1) * n].float().max())
        g = torch.Generator().manual_seed(5)
        per_rank = [torch.randn(world * n, generator=g).to(torch.bfloat16) for _ in range(world)]
        full.copy_(per_rank[rank])
        out = torch.empty(n, device=dev)
        comm.reduce_scatter_avg_into(out, full)
        torch.cuda.synchronize()
        acc = torch.zeros(n)
        for r in range(world):
            acc += per_rank[r][rank * n:(rank + 1) * n].float()
        if comm.nvls_active:       # the switch hands back the bf16-rounded sum
            assert torch.allclose(out.cpu(), acc / world, rtol=2 ** -7, atol=1e-3)
        else:
            assert torch.allclose(out.cpu(), acc / world, atol=1e-6)
        comm.barrier()
        comm.check_errors()
        torch.save({"ok": True, "backend": comm.backend, "multicast": comm.multicast, "nvls": comm.nvls_active},
                   os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        comm.close()
        dist.destroy_process_group()


def watchdog(rank, world, port, out_dir):
    """a peer that never enters a collective must become an error on the ranks that wait for it, within
    the watchdog budget, and the payload must not be touched.  Rank ``world - 1`` skips one all-reduce."""
    import time

    os.environ["DTB_COMM_TIMEOUT_MS"] = "1500"
    from distributed_training_b200.parallel.symm import SymmComm

    _init(rank, world, port)
    comm = SymmComm(torch.device("cuda", rank), segment_bytes=64 << 20)
    try:
        buf = comm.alloc(1 << 20, torch.bfloat16)
        buf.fill_(float(rank + 1))
        comm.all_reduce_avg_(buf).wait()              # everybody takes part: fine
        torch.cuda.synchronize()
        comm.check_errors()
        expect = sum(range(1, world + 1)) / world
        assert float(buf.float().mean()) == expect
        dist.barrier()
        raised = False
        if rank != world - 1:
            buf.fill_(float(rank + 1))
            t0 = time.time()
            comm.all_reduce_avg_(buf).wait()          # the last rank never comes
            torch.cuda.synchronize()
            waited = time.time() - t0
            try:
                comm.check_errors()
            except RuntimeError as exc:
                raised = "never arrived" in str(exc)
            assert raised, "watchdog did not raise"
            assert waited < 20.0, f"watchdog took {waited:.1f} s"
            assert float(buf.float().min()) == float(rank + 1) == float(buf.float().max()), \
                "an abandoned collective must leave the payload alone"
            # later collectives on the dead communicator return at entry instead of spinning again
            t0 = time.time()
            comm.all_reduce_avg_(buf).wait()
            torch.cuda.synchronize()
            assert time.time() - t0 < 1.0
        dist.barrier()
        torch.save({"ok": True, "raised": raised}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        dist.destroy_process_group()


def ddp_native(rank, world, port, out_dir):
    """Native wide-ish MLP under DDPStrategy on W GPUs vs a single-process fp32 autograd reference."""
    import torch.nn.functional as F

    from distributed_training_b200.models.mlp import MLP
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec

    _init(rank, world, port)
    strategy = DDPStrategy(rank, True)
    try:
      for fused in (False, True, "chunked", "adamw", "adamw_chunked"):
        adamw = str(fused).startswith("adamw")   # stateful optimizer: stepped per bucket behind its all-reduce
        if str(fused).endswith("chunked"):   # every parameter in its own bucket + chunked reduction of the last wgrad
            strategy = DDPStrategy(rank, True, comm=strategy.comm, bucket_cap_mb=0.01, first_bucket_mb=0.001)
        torch.manual_seed(100 + rank)                        # rank-dependent init: must be broadcast
        model = MLP(256, 512, 384, num_layers=3, dtype=torch.bfloat16).cuda()
        wrapped = strategy.prepare_model(model, rank)
        assert wrapped.native_step
        if str(fused).endswith("chunked"):
            wrapped.TAIL_CHUNK_MIN_BYTES = 0
            assert wrapped.tail_chunks(model.layers[0].weight) == wrapped.TAIL_CHUNKS
        spec = (OptimizerSpec("adamw", lr=2e-3, weight_decay=0.01) if adamw else
                OptimizerSpec("sgd", lr=0.05, weight_decay=0.01))
        opt = strategy.build_optimizer(wrapped, spec)
        if fused:
            wrapped.attach_optimizer(opt)                    # gradient all-reduce + SGD in one kernel
            assert adamw or wrapped._fused_sgd_hyper() is not None
        torch.manual_seed(100)
        ref = MLP(256, 512, 384, num_layers=3, dtype=torch.bfloat16)
        ref = ref.float().cuda()
        for (k1, v1), (k2, v2) in zip(wrapped.state_dict().items(), ref.state_dict().items()):
            assert k1 == "module." + k2 and torch.equal(v1.float(), v2), k1
        ropt = (torch.optim.AdamW(ref.parameters(), lr=2e-3, weight_decay=0.01) if adamw else
                torch.optim.SGD(ref.parameters(), lr=0.05, weight_decay=0.01))
        g = torch.Generator().manual_seed(9)
        for step in range(4):
            x = torch.randn(world * 128, 256, generator=g).to(torch.bfloat16).cuda()
            y = torch.randint(0, 384, (world * 128,), generator=g).cuda()
            loss = wrapped.forward_backward(x[rank * 128:(rank + 1) * 128], y[rank * 128:(rank + 1) * 128])
            if adamw:    # every parameter was already stepped behind its bucket's all-reduce
                assert not any(opt._pending(grp) for grp in opt.param_groups), "step-behind path did not run"
            opt.step()
            ropt.zero_grad()
            rl = F.cross_entropy(ref(x.float()), y)
            rl.backward()
            ropt.step()
        torch.cuda.synchronize()
        for (k, v), (_, rv) in zip(model.state_dict().items(), ref.state_dict().items()):
            assert torch.allclose(v.float(), rv, atol=3e-2, rtol=3e-2), (fused, k, (v.float() - rv).abs().max())
        flat = torch.cat([p.detach().float().view(-1) for p in model.parameters()])
        got = _gather_cpu(flat, world)
        assert all(torch.equal(got[0], t) for t in got), "replicas diverged"
        # autograd path through the same wrapper (generic models)
        lin = torch.nn.Linear(20, 1).cuda()
        w2 = DDPStrategy(rank, True, comm=strategy.comm).prepare_model(lin, rank)
        o2 = torch.optim.SGD(w2.parameters(), lr=0.1)
        for _ in range(3):
            o2.zero_grad()
            F.mse_loss(w2(torch.randn(32, 20, device="cuda")), torch.randn(32, 1, device="cuda")).backward()
            o2.step()
        got = _gather_cpu(lin.weight.detach(), world)
        assert all(torch.equal(got[0], t) for t in got)
        strategy.comm.check_errors()
        torch.save({"ok": True, "exposed_ms": wrapped.exposed_comm_ms()}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        if strategy.comm is not None:
            strategy.comm.close()
        dist.destroy_process_group()


def sharded_native(rank, world, port, out_dir):
    """FSDPStrategy on the fused kernels (AG->GEMM, RS+optimizer) vs a single-process fp32 reference."""
    import torch.nn.functional as F

    from distributed_training_b200.models.gpt2 import GPT2, GPT2Config
    from distributed_training_b200.models.mlp import MLP
    from distributed_training_b200.parallel.sharded import FSDPStrategy
    from distributed_training_b200.parallel.sharded_native import NativeShardedModule
    from distributed_training_b200.parallel.strategy import OptimizerSpec

    _init(rank, world, port)
    strategy = FSDPStrategy(True)
    try:
        for kind in ("mlp_sgd", "mlp_adamw", "mlp_padded_adamw", "gpt2_adamw"):
            torch.manual_seed(300 + rank)                    # rank-dependent init: must be broadcast
            if kind == "mlp_padded_adamw":
                # 520 rows do not split into equal 16-row-granular shards: the last rank's shard ends in padding
                make = lambda dt: MLP(256, 520, 392, num_layers=3, dtype=dt)
            elif kind.startswith("mlp"):
                make = lambda dt: MLP(256, 512, 384, num_layers=3, dtype=dt)
            else:
                cfg = GPT2Config(vocab_size=1000, n_positions=128, n_embd=128, n_layer=2, n_head=2)
                make = lambda dt: GPT2(cfg, dtype=dt)
            model = make(torch.bfloat16).cuda()
            wrapped = strategy.prepare_model(model, rank)
            assert isinstance(wrapped, NativeShardedModule) and wrapped.native_step
            assert len(wrapped.weights) >= 3
            spec = OptimizerSpec("sgd", lr=0.05) if kind.endswith("sgd") else OptimizerSpec("adamw", lr=2e-3,
                                                                                         weight_decay=0.1)
            opt = strategy.build_optimizer(wrapped, spec)
            torch.manual_seed(300)
            ref = make(torch.bfloat16).float().cuda()        # rank 0's init
            full0 = wrapped.state_dict()
            for k, v in ref.state_dict().items():
                assert torch.equal(full0[k].float(), v.cpu()), k
            ropt = (torch.optim.SGD(ref.parameters(), lr=0.05) if kind.endswith("sgd")
                    else torch.optim.AdamW(ref.parameters(), lr=2e-3, weight_decay=0.1))
            g = torch.Generator().manual_seed(11)
            for step in range(4):
                if kind.startswith("mlp"):
                    x = torch.randn(world * 128, 256, generator=g).to(torch.bfloat16).cuda()
                    y = torch.randint(0, 384, (world * 128,), generator=g).cuda()
                    xs, ys = x[rank * 128:(rank + 1) * 128], y[rank * 128:(rank + 1) * 128]
                    rloss = F.cross_entropy(ref(x.float()), y)
                else:
                    x = torch.randint(0, 1000, (world * 2, 128), generator=g).cuda()
                    y = torch.randint(0, 1000, (world * 2, 128), generator=g).cuda()
                    xs, ys = x[rank * 2:(rank + 1) * 2], y[rank * 2:(rank + 1) * 2]
                    rloss = ref.loss(x, y)
                opt.zero_grad()
                wrapped.forward_backward(xs, ys)
                opt.step()
                ropt.zero_grad()
                rloss.backward()
                ropt.step()
            # ---- FULL_SHARD bookkeeping: full matrices only live in the pools, and the all-gather that runs INSIDE
            # the forward GEMM reproduces every rank's shard bit for bit
            if kind.startswith("mlp"):
                persistent = sum(sw.shard.numel() for sw in wrapped.weights.values())
                total = sum(sw.rows * sw.cols for sw in wrapped.weights.values())
                assert total <= persistent * world <= total + len(wrapped.weights) * 16 * world * 520
                if kind == "mlp_padded_adamw":
                    assert any(sw.padded_rows > sw.rows for sw in wrapped.weights.values())
                assert all(sw.param.data.untyped_storage().nbytes() <= 4 for sw in wrapped.weights.values())
                lin = model.layers[1]
                sw = wrapped.weights[id(lin.weight)]
                if sw.full_slot is not None:
                    sw.full_slot.owner, sw.full_slot = None, None           # force the fused gather
                xin = torch.randn(256, lin.weight.shape[1], device="cuda").to(torch.bfloat16)
                out = torch.empty(256, lin.weight.shape[0], device="cuda", dtype=torch.bfloat16)
                wrapped.gemm_forward(xin, lin.weight, out, m=256, n=lin.weight.shape[0], k=lin.weight.shape[1])
                torch.cuda.synchronize()
                shards = _gather_cpu(sw.shard.contiguous(), world)
                whole = torch.cat(shards, dim=0)[: sw.rows]           # padding rows of the last shard dropped
                assert torch.equal(sw.full.cpu(), whole), "fused gather is not bit-exact"
                ref_out = xin.float() @ whole.float().cuda().t()
                assert torch.allclose(out.float(), ref_out, atol=2e-2, rtol=2e-2)
                # the stand-alone (backward) gather gives the same bytes
                sw.full_slot.owner, sw.full_slot = None, None
                assert torch.equal(wrapped.weight_view(lin.weight).cpu(), whole)
                wrapped._access = []
            full = wrapped.state_dict()                      # collective gather of every shard
            tol = 3e-2 if kind.endswith("sgd") else 2e-2
            for k, v in ref.state_dict().items():
                assert torch.allclose(full[k].float(), v.cpu(), atol=tol, rtol=5e-2), \
                    (kind, k, (full[k].float() - v.cpu()).abs().max())
            got = _gather_cpu(torch.cat([v.float().view(-1) for v in full.values()]).cuda(), world)
            assert all(torch.equal(got[0], t) for t in got), "ranks hold different gathered states"
            # snapshot round trip through the strategy
            path = os.path.join(out_dir, f"snap_{kind}.pt")
            strategy.save_checkpoint(wrapped, {"MODEL_STATE": full, "EPOCHS_RUN": 1}, path)
            dist.barrier()
            wrapped.load_state_dict(torch.load(path, weights_only=False)["MODEL_STATE"])
            strategy.comm.check_errors()
        torch.save({"ok": True}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        if strategy.comm is not None:
            strategy.comm.close()
        dist.destroy_process_group()


def trainer_gpu(rank, world, port, run_dir, total, tag):
    """SURVEY §7.2 minimum slice: Hydra-configured trainer, toy Linear(20,1), DDP strategy on the peer-memory
    kernels (peer-store broadcast, one-shot fused all-reduce from the autograd hook, fused SGD), snapshot, resume."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    os.chdir(run_dir)
    from distributed_training_b200 import hydra_compat as hc
    from distributed_training_b200.ops import native
    from src import distributed_trainer as dt

    snap = os.path.join(run_dir, "snap.pt")
    captured = {}

    def grab(trainer):
        captured["state"] = {k: v.detach().cpu().clone() for k, v in trainer.model.state_dict().items()}
        captured["epochs_run"] = trainer.epochs_run
        captured["launches"] = native.launch_count()
        captured["comm"] = type(trainer._strategy.comm).__name__

    overrides = [f"train.snapshot_path={snap}", "train.dataset_size=512", "train.batch_size=32",
                 "train.dataset=teacher", "train.learning_rate=0.05", f"train.total_epochs={total}",
                 "train.device=cuda", f"hydra.run.dir={run_dir}/run_{tag}"]
    hc.run_job(lambda cfg: dt.run(cfg, before_teardown=grab), ROOT / "conf", "config", overrides)
    torch.save(dict(captured), os.path.join(run_dir, f"res_{tag}_rank{rank}.pt"))


def collectives_stress(rank, world, port, out_dir, max_mb=1024):
    """SURVEY §4.3 / §5.2 stress tier (stands in for the race checks compute-sanitizer would do): every all-reduce
    size class from 1 KiB to ``max_mb`` MiB against an NCCL fp32 oracle, launches queued back to back without host
    synchronisation (flag / sequence reuse across the one-shot and two-shot kernels), and the same collectives while a
    persistent GEMM burst owns the SMs on another stream, with per-rank random host and device launch skew.
    Every result must also be bitwise identical on all ranks."""
    import random
    import time

    from distributed_training_b200.ops import native
    from distributed_training_b200.parallel.symm import SymmComm

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    dev = torch.device("cuda", rank)
    max_bytes = max_mb << 20
    comm = SymmComm(dev, segment_bytes=max_bytes + (192 << 20))
    rng = random.Random(1234 + rank)
    checked = 0

    def fill(n, dtype, seed):
        g = torch.Generator(device=dev).manual_seed(seed * 64 + rank)
        return torch.randn(n, device=dev, generator=g).to(dtype)

    def oracle(x):
        """Mean over ranks by NCCL in fp32 (bf16 inputs are exact in fp32; fp32 sums differ by summation order only)."""
        ref = x.float()
        dist.all_reduce(ref)
        return ref.div_(world)

    def check(got, ref, what):
        nonlocal checked
        tol = 2 ** -7 if got.dtype == torch.bfloat16 else 2e-6
        diff = (got.float() - ref).abs()
        bound = tol * ref.abs() + tol * 4
        assert bool((diff <= bound).all()), (what, float(diff.max()))
        root = got.clone()
        dist.broadcast(root, 0)                     # bitwise identical on every rank
        assert torch.equal(root, got), (what, "ranks disagree bitwise")
        checked += 1

    try:
        big = comm.alloc(max_bytes // 2, torch.bfloat16)
        # ---- A. every size class, idle GPU ----
        sizes = [1 << 10, (4 << 10) + 32, 64 << 10, (64 << 10) + 16, 1 << 20, (16 << 20) + 4096, 128 << 20, max_bytes]
        for nbytes in [s for s in sizes if s <= max_bytes]:
            for dtype in ((torch.bfloat16, torch.float32) if nbytes <= (16 << 20) + 4096 else (torch.bfloat16,)):
                esize = 2 if dtype == torch.bfloat16 else 4
                n = nbytes // esize
                buf = big.view(torch.uint8)[:nbytes].view(dtype)
                for rep in range(2):
                    x = fill(n, dtype, nbytes % 9973 + rep)
                    ref = oracle(x)
                    buf.copy_(x)
                    comm.all_reduce_avg_(buf).wait()
                    check(buf, ref, ("idle", nbytes, str(dtype), rep))
                    del x, ref
        # ---- B. back to back, no host synchronisation, mixed kernels ----
        slot_bytes = [16 << 10, 1 << 20, 48 << 10, (2 << 20) + 64, 1 << 10, 8 << 20, 64 << 10, 512 << 10]
        slots, offs, o = [], [], 0
        raw = big.view(torch.uint8)
        for b in slot_bytes:
            slots.append(raw[o:o + b].view(torch.bfloat16))
            o += (b + 255) // 256 * 256
        for round_ in range(3):
            xs = [fill(s.numel(), torch.bfloat16, 500 + 10 * round_ + i) for i, s in enumerate(slots)]
            refs = [oracle(x) for x in xs]
            for s, x in zip(slots, xs):
                s.copy_(x)
            works = [comm.all_reduce_avg_(s) for s in slots]          # 8 launches queued, none waited for
            works += [comm.all_reduce_avg_(slots[1])]                 # and the mean of a mean on one of them
            for w_ in works:
                w_.wait()
            for i, (s, ref) in enumerate(zip(slots, refs)):
                check(s, ref, ("back-to-back", round_, i))
        # ---- C. beside a GEMM burst that owns the SMs, with launch skew ----
        m, nn, k = 4096, 8192, 8192
        a = fill(m * k, torch.bfloat16, 7).view(m, k)
        bmat = fill(nn * k, torch.bfloat16, 8).view(nn, k)
        out = torch.empty(m, nn, device=dev, dtype=torch.bfloat16)
        gemm_stream = torch.cuda.Stream(dev)
        par = comm.alloc((32 << 20) // 2, torch.bfloat16)
        for it, (nbytes, exclusive) in enumerate([(64 << 10, False), (1 << 20, False), (16 << 20, False),
                                                  (16 << 20, True), (128 << 20, True), (128 << 20, False),
                                                  (32 << 20, True), (32 << 20, False)]):
            if nbytes > max_bytes:
                continue
            n = nbytes // 2
            buf = big[:n]
            x = fill(n, torch.bfloat16, 900 + it)
            ref = oracle(x)
            buf.copy_(x)
            sgd = nbytes == (32 << 20)
            if sgd:
                p0 = fill(n, torch.bfloat16, 77)      # rank 0's draw becomes everybody's parameters
                dist.broadcast(p0, 0)
                par[:n].copy_(p0)
            torch.cuda.synchronize()
            comm.exclusive_now = exclusive
            gemm_ctas = 148 - comm.exclusive_sms if exclusive else 0
            gemm_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(gemm_stream):
                for _ in range(12):                    # ~4 ms of back-to-back persistent GEMMs
                    native.gemm(a, bmat, out, m=m, n=nn, k=k, max_ctas=gemm_ctas)
            time.sleep(rng.uniform(0.0, 2e-3))         # host skew
            torch.cuda._sleep(int(rng.uniform(0, 1.5e6)))     # device skew on the launching stream (up to ~1 ms)
            if sgd:
                comm.all_reduce_sgd_(buf, par[:n], 0.05, 0.01).wait()
            else:
                comm.all_reduce_avg_(buf).wait()
            torch.cuda.current_stream().wait_stream(gemm_stream)
            if sgd:
                expect = p0.float() * (1 - 0.05 * 0.01) - 0.05 * ref
                check(par[:n], expect, ("beside-gemm sgd", nbytes, exclusive))
                assert torch.equal(buf, x), "fused SGD must leave the gradients alone"
            else:
                check(buf, ref, ("beside-gemm", nbytes, exclusive))
        comm.exclusive_now = False
        # all-gather and reduce-scatter beside the same burst
        n = (8 << 20) // 2
        comm.reserve((world + 2) * n * 2 + 4096)
        full, shard = comm.alloc(world * n, torch.bfloat16), comm.alloc(n, torch.bfloat16)
        x = fill(n, torch.bfloat16, 41)
        shard.copy_(x)
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x)
        torch.cuda.synchronize()
        gemm_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(gemm_stream):
            for _ in range(6):
                native.gemm(a, bmat, out, m=m, n=nn, k=k)
        torch.cuda._sleep(int(rng.uniform(0, 1.5e6)))
        comm.all_gather_into(full, shard)
        torch.cuda.synchronize()
        assert torch.equal(full, torch.cat(parts)), "all-gather beside a GEMM"
        xf = fill(world * n, torch.bfloat16, 43)
        ref = oracle(xf)[rank * n:(rank + 1) * n]
        full.copy_(xf)
        rs_out = torch.empty(n, device=dev)
        torch.cuda.synchronize()
        with torch.cuda.stream(gemm_stream):
            for _ in range(6):
                native.gemm(a, bmat, out, m=m, n=nn, k=k)
        time.sleep(rng.uniform(0.0, 2e-3))
        comm.reduce_scatter_avg_into(rs_out, full)
        torch.cuda.synchronize()
        tol = 2 ** -7 if comm.nvls_active else 1e-6
        assert torch.allclose(rs_out, ref, rtol=tol, atol=tol * 4), "reduce-scatter beside a GEMM"
        comm.barrier()
        comm.check_errors()
        torch.save({"ok": True, "checked": checked, "nvls": comm.nvls_active, "max_mb": max_mb},
                   os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        comm.close()
        dist.destroy_process_group()


def ddp_native_gpt2(rank, world, port, out_dir):
    """Native GPT-2 step under DDPStrategy (two-stream backward: weight gradients and every ``mark_ready`` on the side
    stream) against a single-process fp32 autograd model on the global batch: averaged gradients in the buckets, then
    two AdamW steps with identical replicas."""
    from distributed_training_b200.models.gpt2 import GPT2, GPT2Config
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec

    _init(rank, world, port)
    strategy = DDPStrategy(rank, True, bucket_cap_mb=0.25, first_bucket_mb=0.05)     # several parameters per bucket
    try:
        cfg = GPT2Config(vocab_size=1000, n_positions=256, n_embd=128, n_layer=2, n_head=2)
        torch.manual_seed(100 + rank)                 # rank-dependent init: must be broadcast
        model = GPT2(cfg, dtype=torch.bfloat16).cuda()
        wrapped = strategy.prepare_model(model, rank)
        assert wrapped.native_step
        ref = GPT2(cfg, dtype=torch.float32).cuda()
        ref.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
        g = torch.Generator().manual_seed(5)
        tok = torch.randint(0, 1000, (2 * world, 256), generator=g).cuda()
        tgt = torch.randint(0, 1000, (2 * world, 256), generator=g).cuda()
        mine = slice(2 * rank, 2 * rank + 2)
        # ---- gradients: the buckets must hold the mean over ranks = the gradient of the global-batch mean loss
        loss = wrapped.forward_backward(tok[mine], tgt[mine])
        torch.cuda.synchronize()
        ref_loss = ref.loss(tok, tgt)
        ref_loss.backward()
        losses = _gather_cpu(loss.detach().float().view(1), world)
        assert abs(float(sum(losses)) / world - float(ref_loss)) < 5e-2
        bad = []
        for (name, p), (_, rp) in zip(model.named_parameters(), ref.named_parameters()):
            got = wrapped.grad_view(p).float().view_as(rp.grad)
            scale = rp.grad.abs().max().item() + 1e-6
            err = (got - rp.grad).abs().max().item()
            if err > 0.08 * scale + 2e-4:
                bad.append((name, err, scale))
        assert not bad, bad[:6]
        # ---- two optimizer steps (AdamW stepped per bucket behind its all-reduce): replicas stay identical
        opt = strategy.build_optimizer(wrapped, OptimizerSpec("adamw", lr=1e-3, weight_decay=0.01))
        wrapped.attach_optimizer(opt)
        for step in range(2):
            wrapped.forward_backward(tok[mine], tgt[mine])
            opt.step()
        torch.cuda.synchronize()
        flat = torch.cat([p.detach().float().view(-1) for p in model.parameters()])
        got = _gather_cpu(flat, world)
        assert all(torch.equal(got[0], t) for t in got), "replicas diverged"
        assert bool(torch.isfinite(flat).all())
        strategy.comm.check_errors()
        torch.save({"ok": True}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        if strategy.comm is not None:
            strategy.comm.close()
        dist.destroy_process_group()
