"""CPU: the C-ABI library loads and exports every symbol include/b2d.h declares (no compute calls),
and the host-side logic (weight packing, gradient layout, state_dict keys, error behaviour)."""
import copy
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "b2d.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2d_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    from lipreading_video_generation_b200 import _lib
    from lipreading_video_generation_b200.build import build
    build()
    lib = _lib.lib()
    names = _declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b2d.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    for n in _lib.SIGNATURES:
        assert n in names, f"{n} bound in _lib.py but not declared in include/b2d.h"
    assert lib.b2d_version() >= 100


def test_no_cpu_fallback_cpu_tensors_raise():
    import lipreading_video_generation_b200 as P
    s = P.LinearNoiseScheduler(10, 1e-4, 1e-2)
    x = torch.zeros(2, 3, 4, 4)
    with pytest.raises(RuntimeError):
        s.add_noise(x, x, torch.zeros(2, dtype=torch.long))
    with pytest.raises(RuntimeError):
        P.MSELoss()(x, x)
    m = P.UNetModel(16, 64, 64, 3, 1, (1,), channel_mult=(1,))
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 64, 16, 16), torch.zeros(1, dtype=torch.long))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "lipreading_video_generation_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert "oracle" not in re.sub(r'""".*?"""', "", src, flags=re.S), f


def test_pack_conv_weight_layout():
    from lipreading_video_generation_b200 import ops
    w = torch.arange(2 * 3 * 3 * 3, dtype=torch.float32).reshape(2, 3, 3, 3)
    p = ops.pack_conv_weight(w, cin_pad=4)
    assert p.shape == (16, 9 * 4) and p.dtype == torch.float16
    p3 = p.view(16, 9, 4).float()
    for co in range(2):
        for ky in range(3):
            for kx in range(3):
                for ci in range(3):
                    assert p3[co, ky * 3 + kx, ci] == w[co, ci, ky, kx]
    assert p3[2:].abs().sum() == 0 and p3[:, :, 3].abs().sum() == 0


def test_pack_dgrad_weight_is_flipped_transpose():
    """dX = conv(dY, flip(w)^T): check against autograd on CPU with exactly representable values."""
    from lipreading_video_generation_b200 import ops
    g = torch.Generator().manual_seed(0)
    w = torch.randint(-3, 4, (16, 16, 3, 3), generator=g).float()
    p = ops.pack_dgrad_weight(w).float().view(16, 9, 16)              # [ci][tap][co]
    w_d = p.view(16, 3, 3, 16).permute(0, 3, 1, 2).contiguous()        # OIHW of the dgrad convolution
    x = torch.randint(-2, 3, (1, 16, 5, 5), generator=g).float().requires_grad_(True)
    y = torch.nn.functional.conv2d(x, w, padding=1)
    dy = torch.randint(-2, 3, y.shape, generator=g).float()
    y.backward(dy)
    assert torch.equal(x.grad, torch.nn.functional.conv2d(dy, w_d, padding=1))


def _tiny(P, **kw):
    class Stub(torch.nn.Module):
        def forward(self, a):
            return a["feats"]
    cfg = dict(image_size=32, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=1,
               attention_resolutions=(1, 2), channel_mult=(1, 2), dropout=0.0, audio_feature_dim=768,
               projected_audio_dim=128, audio_encoder=Stub())
    cfg.update(kw)
    return P.UNetAudio(**cfg)


def test_grad_layout_covers_every_torso_parameter_once_and_unpacks_to_param_shape():
    import lipreading_video_generation_b200 as P
    from lipreading_video_generation_b200.engine import _Grads
    m = _tiny(P)
    eng = m.engine()
    params = eng.torso_params()
    ids = [id(p) for p in params]
    assert len(ids) == len(set(ids))
    assert set(ids) == {id(p) for p in m.parameters()}         # stub encoder has no parameters
    G = _Grads(eng, "cpu")
    G.flat.copy_(torch.arange(G.flat.numel(), dtype=torch.float32))
    for p in params:
        assert tuple(G.unpacked(p).shape) == tuple(p.shape)
    # OIHW view of a packed [co][tap][ci] buffer
    conv = m.input_blocks[1][0].in_layers[2]
    buf, v = G.g(conv.weight), G.unpacked(conv.weight)
    assert v[5, 7, 2, 1] == buf.view(buf.shape[0], 9, -1)[5, 2 * 3 + 1, 7]
    # stage offsets are monotone and end at the buffer size
    assert G.stage_end == sorted(G.stage_end) and G.stage_end[-1] == G.flat.numel()


def test_state_dict_keys_match_reference_layout(golden):
    """Drop-in: the keys of the reference checkpoint load unchanged."""
    import lipreading_video_generation_b200 as P
    m = _tiny(P)
    keys = set(m.state_dict().keys())
    g = golden("unet_tiny.pt")
    assert set(g["grad_norms"].keys()) == keys                  # keys recorded from the real reference model
    for k in ("time_embed.0.weight", "input_blocks.0.0.weight", "input_blocks.1.0.in_layers.2.weight",
              "input_blocks.1.0.emb_layers.1.bias", "input_blocks.1.1.qkv.weight", "input_blocks.2.0.op.weight",
              "middle_block.1.proj_out.weight", "output_blocks.1.2.conv.weight", "out.2.bias",
              "audio_transformer.transform.0.weight", "cond_conv_in.weight"):
        assert k in keys, k


@pytest.mark.skipif(not __import__("oracle.ref_import", fromlist=["x"]).available(), reason="/root/reference absent")
def test_full_config_state_dict_equals_live_reference():
    """train.py:88-97 configuration: same 309 non-wav2vec2 keys and shapes as the real reference class."""
    import lipreading_video_generation_b200 as P
    from oracle import ref_import
    ref = ref_import.load()

    class Stub(torch.nn.Module):
        def forward(self, a):
            return a["feats"]
    kw = dict(image_size=128, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
              attention_resolutions=(1, 2, 4), audio_feature_dim=768, projected_audio_dim=128)
    mine = P.UNetAudio(**kw, audio_encoder=Stub())
    rm = ref.unet.UNetModel(128, 195, 64, 3, 2, (1, 2, 4), dropout=0.1, channel_mult=(1, 2, 4))
    rsd = {k: tuple(v.shape) for k, v in rm.state_dict().items()}
    msd = {k: tuple(v.shape) for k, v in mine.state_dict().items()
           if not k.startswith(("audio_transformer", "cond_conv_in"))}
    assert rsd == msd
    assert sum(p.numel() for p in rm.parameters()) == 16250883
    assert len(mine.state_dict()) == 309
    # zero_module'd tensors are zero at init, as in the reference
    for k, v in mine.state_dict().items():
        if ".out_layers.3." in k or ".proj_out." in k or k.startswith("out.2."):
            assert v.abs().sum() == 0, k


def test_unsupported_constructor_options_raise_not_implemented():
    import lipreading_video_generation_b200 as P
    for kw in (dict(dims=1), dict(num_classes=10), dict(use_scale_shift_norm=True), dict(resblock_updown=True),
               dict(use_fp16=True), dict(conv_resample=False)):
        with pytest.raises(NotImplementedError):
            P.UNetModel(32, 64, 64, 3, 1, (1,), channel_mult=(1,), **kw)


def test_sampler_index_errors_match_reference_behaviour():
    import lipreading_video_generation_b200 as P
    s = P.LinearNoiseScheduler(100, 0.00085, 0.012)
    x = torch.zeros(1, 3, 4, 4)
    with pytest.raises(IndexError):            # train.py:125 draws t in [0,500) for a 100-step table
        s.sample_prev_timestep(x, x, torch.tensor([499]))
    with pytest.raises(RuntimeError):          # a (B,) t with B > 1 does not broadcast in the reference either
        s.sample_prev_timestep(x, x, torch.tensor([1, 2]))


# ------------------------------------------------------------------ wav2vec2 positional conv, short sequences
@pytest.mark.parametrize("T", [1, 12, 32, 40])
def test_pos_conv_short_sequence_matches_stock_module(T):
    """unet_audio._pos_conv_short_sequence restates HF's Wav2Vec2PositionalConvEmbedding.forward as one batched
    GEMM per group for T <= 32 frames (train.py feeds 12): same outputs and gradients as the stock module,
    fp32 tolerance 1e-5 relative; T = 40 takes the stock path."""
    import types
    from transformers import Wav2Vec2Config
    from transformers.models.wav2vec2.modeling_wav2vec2 import Wav2Vec2PositionalConvEmbedding
    from lipreading_video_generation_b200 import unet_audio as UA

    torch.manual_seed(0)
    cfg = Wav2Vec2Config(hidden_size=64, num_conv_pos_embeddings=128, num_conv_pos_embedding_groups=16)
    stock = Wav2Vec2PositionalConvEmbedding(cfg)
    fast = copy.deepcopy(stock)
    fast._b2d_stock_forward = fast.forward
    fast.forward = types.MethodType(UA._pos_conv_short_sequence, fast)
    x = torch.randn(5, T, 64)
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    y1, y2 = stock(x1), fast(x2)
    assert y1.shape == y2.shape == (5, T, 64)
    r = torch.randn_like(y1)
    (y1 * r).sum().backward()
    (y2 * r).sum().backward()

    def rel(a, b):
        return float((a - b).norm() / b.norm())
    assert rel(y2, y1) <= 1e-5
    assert rel(x2.grad, x1.grad) <= 1e-5
    for (n1, p1), (_, p2) in zip(stock.named_parameters(), fast.named_parameters()):
        # the weight-norm w = g v / ||v|| is evaluated with plain tensor ops (unet_audio._weight_normed) instead of
        # torch._weight_norm: same formula, another fp32 summation order over the 64 x 4 x 128 weights
        assert rel(p2.grad, p1.grad) <= 1e-4, n1


def test_single_channel_conv1d_matches_stock_module():
    """unet_audio._conv1d_single_channel (the waveform conv of the wav2vec2 feature extractor as a window view
    times the weight) against nn.Conv1d: outputs and gradients, fp32 tolerance 1e-5 relative."""
    import types
    from lipreading_video_generation_b200 import unet_audio as UA

    torch.manual_seed(0)
    for bias in (False, True):
        stock = torch.nn.Conv1d(1, 32, kernel_size=10, stride=5, bias=bias)
        fast = copy.deepcopy(stock)
        fast._b2d_stock_forward = fast.forward
        fast.forward = types.MethodType(UA._conv1d_single_channel, fast)
        x = torch.randn(4, 1, 403)
        y1, y2 = stock(x), fast(x)
        assert y1.shape == y2.shape == (4, 32, 79)
        r = torch.randn_like(y1)
        (y1 * r).sum().backward()
        (y2 * r).sum().backward()
        assert float((y2 - y1).norm() / y1.norm()) <= 1e-5
        for (n1, p1), (_, p2) in zip(stock.named_parameters(), fast.named_parameters()):
            assert float((p2.grad - p1.grad).norm() / p1.grad.norm()) <= 1e-5, n1


def test_dims3_module_tree_has_the_reference_keys_and_conv3d_shapes():
    """UNetModel(dims=3): same state_dict keys as dims=2, Conv3d (O, I, 3, 3, 3) weights,
    Downsample stride (1, 2, 2) (unet.py:141); where /root/reference is mounted, keys and shapes equal the live class."""
    import lipreading_video_generation_b200 as P
    kw = dict(image_size=16, in_channels=64, model_channels=64, out_channels=3, num_res_blocks=1,
              attention_resolutions=(1, 2), channel_mult=(1, 2))
    m3, m2 = P.UNetModel(dims=3, **kw), P.UNetModel(dims=2, **kw)
    assert list(m3.state_dict()) == list(m2.state_dict())
    assert tuple(m3.input_blocks[0][0].weight.shape) == (64, 64, 3, 3, 3)
    assert tuple(m3.input_blocks[2][0].op.stride) == (1, 2, 2)
    assert tuple(m3.input_blocks[3][0].skip_connection.weight.shape) == (128, 64, 1, 1, 1)
    if os.path.isdir("/root/reference/video-generation/diffusion"):
        from oracle import ref_import
        ref = ref_import.load()
        r = ref.unet.UNetModel(dims=3, **kw)
        assert {k: tuple(v.shape) for k, v in r.state_dict().items()} == \
               {k: tuple(v.shape) for k, v in m3.state_dict().items()}


def test_feature_layer0_patch_is_installed_and_cpu_tensors_take_the_stock_path():
    """patch_feature_layer0 only reroutes CUDA tensors (the kernel is the product path there); on CPU tensors the
    third-party layer runs untouched, so the patched module equals the stock one bit for bit."""
    import copy
    from transformers import Wav2Vec2Config, Wav2Vec2Model
    import lipreading_video_generation_b200.unet_audio as UA
    torch.manual_seed(0)
    m = Wav2Vec2Model(Wav2Vec2Config(num_hidden_layers=1))
    p = UA.patch_feature_layer0(copy.deepcopy(m))
    l0 = p.feature_extractor.conv_layers[0]
    assert hasattr(l0, "_b2d_stock_forward")
    a = torch.randn(2, 1, 4000)
    assert torch.equal(l0(a), m.feature_extractor.conv_layers[0](a))
    UA.patch_feature_layer0(p)                                   # idempotent
    assert l0._b2d_stock_forward.__func__ is type(l0).forward


def test_short_attention_interface_falls_back_to_library_sdpa_on_cpu_tensors():
    """patch_short_attention registers the kernel through transformers' attention-interface registry; CPU tensors
    (and any shape the kernel is not built for) are served by the library SDPA function, so the patched encoder
    equals the stock one on CPU."""
    import copy
    from transformers import Wav2Vec2Config, Wav2Vec2Model
    import lipreading_video_generation_b200.unet_audio as UA
    torch.manual_seed(0)
    stock = Wav2Vec2Model(Wav2Vec2Config(num_hidden_layers=2)).eval()
    fast = UA.patch_short_attention(copy.deepcopy(stock))
    assert fast.config._attn_implementation == UA.SHORT_ATTENTION_NAME
    a = torch.randn(2, 4000)
    n0 = UA._short_attention_calls[0]
    with torch.no_grad():
        assert torch.equal(fast(input_values=a).last_hidden_state, stock(input_values=a).last_hidden_state)
    assert UA._short_attention_calls[0] == n0


def test_stock_route_counter_counts_cpu_hand_offs_and_strict_mode_leaves_cpu_tensors_alone(monkeypatch):
    """Every patched encoder piece counts the calls it hands to the stock library path ("at least count and
    expose the fallbacks").  On CPU tensors everything is a hand-off: the counter must name the patch sites, reset on
    request, and B2D_W2V_STRICT=1 must not raise for CPU tensors (it guards the GPU product path only).  The cast cache
    of Wav2Vec2Encoder stays off outside CUDA + autocast + training."""
    import lipreading_video_generation_b200.unet_audio as UA
    monkeypatch.setenv("B2D_W2V_STRICT", "1")
    torch.manual_seed(0)
    enc = UA.Wav2Vec2Encoder(pretrained=False).eval()
    UA.stock_route_counts(reset=True)
    with torch.no_grad():
        out = enc({"input_values": torch.randn(2, 4000)})
    assert out.shape == (2, 12, 768)
    routes = UA.stock_route_counts()
    for site in ("feature_layer0", "layer_norm", "short_attention", "feature_convs", "linear_eval_or_fp32"):
        assert routes.get(site, 0) > 0, (site, routes)
    assert routes["short_attention"] == 12 and routes["feature_convs"] == 1 and routes["feature_layer0"] == 1
    assert UA.stock_route_counts(reset=True) == routes and UA.stock_route_counts() == {}
    assert enc._cast_cache is None and UA._CAST_CACHE_ACTIVE == [False]


def test_every_environment_switch_the_code_reads_is_documented():
    """Each B2D_* variable read with getenv / os.environ in the package, the bench or the entry points appears in
    INTEGRATION.md's switch table (or DESIGN.md): an undocumented switch is an undocumented code path."""
    import glob
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "lipreading_video_generation_b200")
    files = (glob.glob(os.path.join(pkg, "*.py")) + glob.glob(os.path.join(pkg, "csrc", "*.cu")) +
             glob.glob(os.path.join(pkg, "csrc", "*.cuh")) + [os.path.join(root, "bench.py"), os.path.join(root, "__graft_entry__.py")])
    pat = re.compile(r'(?:getenv\(\s*|environ\.get\(\s*|environ\[\s*|setdefault\(\s*)"(B2D_[A-Z0-9_]+)"')
    read = set()
    for f in files:
        read |= set(pat.findall(open(f).read()))
    assert len(read) >= 20, read                      # the pattern still matches how the code reads its switches
    docs = open(os.path.join(root, "INTEGRATION.md")).read() + open(os.path.join(root, "DESIGN.md")).read()
    missing = sorted(v for v in read if v not in docs)
    assert not missing, missing


# ----------------------------------------------------------------------------- Adam (row N3) host logic
def test_adam_chunk_table_splits_tensors_into_runs_and_matches_the_c_struct():
    import ctypes as C
    from lipreading_video_generation_b200 import optim
    assert optim._CHUNK_DTYPE.itemsize == 40          # sizeof(B2dAdamChunk): 4 pointers + int64
    t = optim.build_chunk_table([(1000, 2000, 3000, 4000, 10), (1 << 20, 2 << 20, 3 << 20, 4 << 20, 2 * 65536 + 5)], chunk=65536)
    assert len(t) == 4
    assert t["n"].tolist() == [10, 65536, 65536, 5]
    assert t["param"].tolist() == [1000, 1 << 20, (1 << 20) + 4 * 65536, (1 << 20) + 8 * 65536]
    assert t["exp_avg_sq"][3] == (4 << 20) + 8 * 65536
    assert sum(t["n"].tolist()) == 10 + 2 * 65536 + 5
    assert len(optim.build_chunk_table([(8, 8, 8, 8, 0)])) == 0       # empty tensor: no chunk


def test_adam_rejects_cpu_parameters_and_unused_options():
    import lipreading_video_generation_b200 as P
    w = torch.nn.Parameter(torch.ones(4))
    w.grad = torch.ones(4)
    with pytest.raises(RuntimeError):
        P.Adam([w], lr=1e-2).step()
    with pytest.raises(NotImplementedError):
        P.Adam([w], amsgrad=True)
    with pytest.raises(ValueError):
        P.Adam([w], lr=-1.0)
    o = P.Adam([w], lr=1e-2)
    assert set(o.state_dict()["param_groups"][0]) >= {"lr", "betas", "eps", "weight_decay"}


def test_pack_table_block_ranges_are_contiguous_and_match_the_c_struct():
    from lipreading_video_generation_b200 import ops
    assert ops._PACK_DESC.itemsize == 72             # sizeof(B2dPackDesc)
    recs = [(4096, 8192, 576, 9, 1, 64, 64, 9, 0, 64, 64),       # 64x64 3x3 fprop: 36864 elements -> 9 CTAs of 4096
            (4096, 16384, 576, 9, 1, 64, 64, 9, 1, 64, 64),      # its dgrad copy
            (1 << 20, 1 << 21, 195, 1, 1, 3, 195, 1, 0, 16, 256)]  # padded 1x1: 4096 elements -> 1 CTA
    t, blocks = ops.build_pack_table(recs, 4096)
    assert t["block_begin"].tolist() == [0, 9, 18] and blocks == 19
    assert t["out"].tolist() == [8192, 16384, 1 << 21] and t["transpose"].tolist() == [0, 1, 0]
