"""CPU, world_size 2 over gloo: the data-parallel host logic (bucket planning, stage-triggered all-reduce of the
engine's flat gradient buffer, parameter broadcast, clip sharding).  The NCCL path itself runs on the GPU box."""
import os
import socket
import sys

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_clips():
    from lipreading_video_generation_b200.parallel import shard_range
    for n, w in ((16, 8), (17, 4), (3, 8), (0, 2)):
        spans = [shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        for a, b in zip(spans, spans[1:]):
            assert a[1] == b[0]
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1


def test_plan_buckets_covers_buffer_in_stage_order():
    from lipreading_video_generation_b200.parallel import plan_buckets
    ends = [10, 10, 50, 120, 121, 300]
    b = plan_buckets(ends, 100)
    assert b[0][1] == 0 and b[-1][2] == 300
    for x, y in zip(b, b[1:]):
        assert x[2] == y[1] and x[0] < y[0]
    assert all(hi - lo >= 100 for _, lo, hi in b[:-1])
    assert plan_buckets([5], 100) == [(0, 0, 5)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), B2D_WAV2VEC2_RANDOM_INIT="1")
        import torch.distributed as dist
        dist.init_process_group("gloo", rank=rank, world_size=world)
        import lipreading_video_generation_b200 as P
        from lipreading_video_generation_b200.engine import _Grads
        from lipreading_video_generation_b200.parallel import GradientAllReducer, broadcast_parameters

        class Enc(torch.nn.Module):                  # stands for the wav2vec2 encoder: a non-torso parameter
            def __init__(self):
                super().__init__()
                self.w = torch.nn.Parameter(torch.zeros(7))
                self.dropped = torch.nn.Parameter(torch.zeros(5))     # a layer LayerDrop skips on rank 1 only

            def forward(self, a):
                return a["feats"]
        torch.manual_seed(100 + rank)               # different init per rank: broadcast must equalise it
        m = P.UNetAudio(image_size=32, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=1,
                        attention_resolutions=(1, 2), channel_mult=(1, 2), dropout=0.0, audio_feature_dim=768,
                        projected_audio_dim=128, audio_encoder=Enc())
        broadcast_parameters(m)
        chk = torch.stack([p.detach().double().sum() for p in m.parameters()])
        gathered = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(gathered, chk)
        assert all(torch.equal(gathered[0], g) for g in gathered), "parameters differ after broadcast"

        red = GradientAllReducer(m, bucket_mb=0.5)
        eng = m.engine()
        G = _Grads(eng, "cpu")
        n = G.flat.numel()
        base = torch.arange(n, dtype=torch.float32) % 97
        G.flat.copy_(base * (rank + 1))
        # replay the backward: stage k finishes -> hook; a bucket is reduced only once all its stages are done
        reduced_upto = 0
        for stage in range(eng.n_stages):
            m._grad_stage_hook(G, stage)
            done = [b for b in red._buckets if b[0] <= stage]
            reduced_upto = done[-1][2] if done else 0
            want_done = base[:reduced_upto] * (sum(range(1, world + 1)) / world)
            assert torch.allclose(G.flat[:reduced_upto], want_done), f"stage {stage}"
            assert torch.equal(G.flat[reduced_upto:], base[reduced_upto:] * (rank + 1)), "bucket reduced too early"
        m._grad_final_hook(G)
        assert reduced_upto == n
        assert len(red._buckets) > 1
        # non-torso parameters: fixed flat layout over ALL of them; a gradient missing on one rank counts as zeros
        # (round-1 advisor finding: ranks all-reduced flat tensors of different lengths when LayerDrop differed)
        m.audio_encoder.w.grad = torch.full((7,), float(rank + 1))
        if rank == 0:
            m.audio_encoder.dropped.grad = torch.full((5,), 4.0)
        red.finish()
        assert torch.allclose(m.audio_encoder.w.grad, torch.full((7,), sum(range(1, world + 1)) / world))
        assert torch.allclose(m.audio_encoder.dropped.grad, torch.full((5,), 4.0 / world))
        assert m.audio_encoder.w.grad.data_ptr() == red._views[red._index[id(m.audio_encoder.w)]].data_ptr()
        assert red.bytes_reduced == 4 * (n + 8 + 8)               # views are padded to 16 bytes
        # the hook path: gradients arriving through autograd trigger their bucket without finish() copying back
        m.audio_encoder.w.grad = None
        m.audio_encoder.dropped.grad = None
        (m.audio_encoder.w * float(rank + 1)).sum().backward()
        red.finish()
        assert torch.allclose(m.audio_encoder.w.grad, torch.full((7,), sum(range(1, world + 1)) / world))
        assert float(m.audio_encoder.dropped.grad.abs().sum()) == 0.0
        # the engine's end-of-backward hook issues the remaining buckets; finish() must not issue them a second time
        # (the 378 MB of encoder gradients were all-reduced twice per step before this was checked)
        (m.audio_encoder.w * float(rank + 1)).sum().backward()
        b0 = red.bytes_reduced
        m._grad_backward_end_hook()
        red.finish()
        assert red.bytes_reduced - b0 == 4 * (8 + 8)
        # no-overlap mode: nothing is reduced while the stages finish; one collective per flat buffer at the end
        red.detach()
        red2 = GradientAllReducer(m, bucket_mb=0.5, overlap=False)
        G.flat.copy_(base * (rank + 1))
        for stage in range(eng.n_stages - 1):
            m._grad_stage_hook(G, stage)
            assert torch.equal(G.flat, base * (rank + 1)), "reduced before the last stage in no-overlap mode"
        m._grad_stage_hook(G, eng.n_stages - 1)
        m._grad_final_hook(G)
        assert len(red2._buckets) == 1
        assert torch.allclose(G.flat, base * (sum(range(1, world + 1)) / world))
        m.audio_encoder.w.grad = None
        m.audio_encoder.dropped.grad = None
        (m.audio_encoder.w * float(rank + 1)).sum().backward()
        assert red2.bytes_reduced == 4 * n, "a gradient hook issued a collective in no-overlap mode"
        red2.finish()
        assert red2.bytes_reduced == 4 * (n + 8 + 8)
        assert torch.allclose(m.audio_encoder.w.grad, torch.full((7,), sum(range(1, world + 1)) / world))
        assert float(m.audio_encoder.dropped.grad.abs().sum()) == 0.0
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc() + repr(e)))


def test_gradient_all_reducer_world2_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res


# ----------------------------------------------------------------------------- sequence-parallel attention (row N1)
def test_shard_rows_gives_equal_aligned_query_shares():
    from lipreading_video_generation_b200.parallel import shard_rows
    assert [shard_rows(1048576, r, 8) for r in (0, 7)] == [(0, 131072), (917504, 131072)]   # cfg5, 8 GPUs
    assert [shard_rows(4096, r, 2) for r in range(2)] == [(0, 2048), (2048, 2048)]
    assert shard_rows(4096, 0, 1) is None                  # single rank: replicated path
    assert shard_rows(1024, 0, 8) is None                  # 128 rows per rank: not 256-aligned -> layer runs replicated
    assert shard_rows(768, 0, 2) is None


def _sp_worker(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        import torch.distributed as dist
        dist.init_process_group("gloo", rank=rank, world_size=world)
        import lipreading_video_generation_b200 as P
        from lipreading_video_generation_b200.parallel import SequenceParallelAttention
        m = P.UNetModel(32, 64, 64, 3, 1, (1,), channel_mult=(1,), dims=3)
        sp = SequenceParallelAttention(m, min_rows=256)
        assert m.engine().seq_parallel is sp
        L_, C_ = 1024, 8
        assert sp.rows(1, L_, 64) == (rank * 512, 512) and sp.rows(1, L_, 128) == (rank * 512, 512)
        assert sp.rows(2, L_, 64) is None                  # two samples: shares are not contiguous
        assert sp.rows(1, L_, 256) is None                 # no row-range forward kernel for head dim 256
        assert sp.rows(1, 256, 64) is None                 # too short to be worth two collectives
        # forward contract: each rank fills ITS rows of `out`, the gather completes it in place
        full = torch.arange(L_ * C_, dtype=torch.float32).view(1, 32, 32, C_)
        out = torch.full_like(full, -1.0)
        r0, n = sp.rows(1, L_, 64)
        out.view(-1, C_)[r0:r0 + n] = full.view(-1, C_)[r0:r0 + n]
        sp.all_gather_rows(out, (r0, n))
        assert torch.equal(out, full)
        lse = torch.full((1, 1, L_), -1.0)
        lse.view(-1)[r0:r0 + n] = torch.arange(r0, r0 + n, dtype=torch.float32)
        sp.all_gather_lse(lse, (r0, n))
        assert torch.equal(lse.view(-1), torch.arange(L_, dtype=torch.float32))
        # backward contract: dQ rows only from their owner (zero elsewhere), dK / dV partial everywhere -> one sum
        dqkv = torch.zeros(1, 32, 32, 3 * C_)
        dqkv.view(-1, 3 * C_)[r0:r0 + n, :C_] = 5.0
        dqkv.view(-1, 3 * C_)[:, C_:] = float(rank + 1)
        sp.all_reduce_sum(dqkv)
        assert (dqkv.view(-1, 3 * C_)[:, :C_] == 5.0).all() and (dqkv.view(-1, 3 * C_)[:, C_:] == 3.0).all()
        assert sp.bytes_moved > 0
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc() + repr(e)))


def test_sequence_parallel_attention_contract_world2_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sp_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res
