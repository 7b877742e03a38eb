"""Host-side logic that needs no GPU: the C-ABI library loads and exports every symbol of include/dtb200.h, parameter layouts match the
oracle's state_dict, the product refuses to compute on CPU (no fallback), and the data-parallel plumbing works at world_size 2 on gloo."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from mtrl_b200 import _lib

    lib = _lib.load()
    syms = _lib.exported_symbols()
    assert len(syms) >= 30
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    assert lib.dtb_abi_version() == 1


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "multimodal-transformer-rl_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "/root/reference" not in txt, f


@pytest.mark.parametrize("cfg", [dict(state_dim=11, act_dim=3, hidden_size=128, n_layer=3, n_head=1),
                                 dict(state_dim=1024, act_dim=18, hidden_size=256, n_layer=2, n_head=8, action_tanh=False),
                                 dict(state_dim=11, act_dim=3, hidden_size=768, n_layer=2, n_head=12, activation_function="gelu_new")])
def test_dt_layout_and_state_dict_roundtrip(cfg):
    from mtrl_b200.model import DecisionTransformer
    from oracle.decision_transformer import DecisionTransformer as Oracle

    o = Oracle(max_length=20, max_ep_len=100, **cfg)
    m = DecisionTransformer(max_length=20, max_ep_len=100, device="cpu", **cfg)
    sd_o, sd_m = o.state_dict(), m.state_dict()
    assert set(sd_o) == set(sd_m)
    assert all(sd_o[k].shape == sd_m[k].shape for k in sd_o)
    m.load_state_dict(sd_o)
    assert all(torch.equal(m.state_dict()[k], sd_o[k]) for k in sd_o)
    # every parameter is a view into the single flat buffer, 8-element aligned, stages cover it exactly once
    base = m.flat_params.data_ptr()
    for n, p in m.named_parameters():
        off = (p.data_ptr() - base) // 4
        assert 0 <= off < m._total and off % 8 == 0, n
    ranges = sorted(m.stage_grad_range(s) for s in range(m.n_stages))
    assert ranges[0][0] == 0 and ranges[-1][1] == m._total
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    o2 = Oracle(max_length=20, max_ep_len=100, **cfg)
    o2.load_state_dict(m.state_dict())  # and back
    # HF-prefixed keys are accepted
    hf_style = {("encoder." + k[len("transformer."):] if k.startswith("transformer.") else k): v for k, v in sd_o.items()}
    hf_style["encoder.wpe.weight"] = torch.zeros(4, cfg["hidden_size"])
    m.load_state_dict(hf_style)


def test_encoder_layout_and_state_dict_roundtrip():
    from mtrl_b200.encoder import MultimodalStateEncoder
    from oracle.encoders import MultimodalStateEncoder as Oracle

    for args in [((1, 84, 84), (1, 84, 84), "big", "concat"), ((3, 64, 64), (2, 64, 64), "small", "sum")]:
        o, e = Oracle(*args), MultimodalStateEncoder(*args, device="cpu")
        assert set(o.state_dict()) == set(e.state_dict()) and o.state_dim == e.state_dim
        e.load_state_dict(o.state_dict())
        assert all(torch.equal(e.state_dict()[k], v) for k, v in o.state_dict().items())
        # flat layout for the data-parallel exchange: both towers' conv tensors first, both FC tensors last (one contiguous FC range), the
        # entries never overlap, and the names keep the reference's module order
        fc_begin, fc_end = e.fc_grad_range()
        assert 0 < fc_begin < fc_end == e._total
        fc_idx = "7" if args[2] == "big" else "5"
        spans = []
        for name, shp, off in zip(e._names, e._shapes, e._offsets):
            n = int(torch.Size(shp).numel())
            is_fc = name.split(".")[1] == fc_idx
            assert (off >= fc_begin) == is_fc, (name, off, fc_begin)
            spans.append((off, off + n))
        spans.sort()
        assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))
        assert [n.split(".")[0] for n in e._names[:2]] == ["video_net", "video_net"] and e._names[-1].startswith("audio_net")


def test_no_cpu_fallback():
    from mtrl_b200 import _lib
    from mtrl_b200.encoder import MultimodalStateEncoder
    from mtrl_b200.model import DecisionTransformer

    m = DecisionTransformer(11, 3, 128, max_length=5, max_ep_len=50, device="cpu")
    s, a, r, t = torch.zeros(1, 5, 11), torch.zeros(1, 5, 3), torch.zeros(1, 5, 1), torch.zeros(1, 5, dtype=torch.long)
    with pytest.raises(_lib.DtbError):
        m(s, a, None, r, t)
    with pytest.raises(_lib.DtbError):
        m.get_action(s[0], a[0], None, r[0], t[0])
    e = MultimodalStateEncoder((1, 84, 84), (1, 84, 84), device="cpu")
    with pytest.raises(_lib.DtbError):
        e(torch.zeros(1, 1, 1, 84, 84), torch.zeros(1, 1, 1, 84, 84))


def test_bad_configs_are_rejected():
    from mtrl_b200 import _lib
    from mtrl_b200.model import DecisionTransformer

    with pytest.raises(_lib.DtbError):
        DecisionTransformer(11, 3, 100, n_head=1, device="cpu")  # head_dim 100 unsupported
    with pytest.raises(ValueError):
        DecisionTransformer(11, 3, 128, activation_function="swish", device="cpu")


def test_shard_batch():
    from mtrl_b200.dp import shard_batch

    x, y = torch.arange(24).reshape(8, 3), torch.arange(8)
    parts = [shard_batch([x, y], r, 4) for r in range(4)]
    assert torch.equal(torch.cat([p[0] for p in parts]), x) and torch.equal(torch.cat([p[1] for p in parts]), y)


_DP_SCRIPT = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from mtrl_b200.dp import GradAllReducer, broadcast_parameters, replicas_consistent
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
torch.manual_seed(rank)
params = torch.randn(1000)
broadcast_parameters([params])
assert replicas_consistent([params])
# emulate a staged backward: ranges become ready from the tail downwards; local grads are pre-scaled by 1/world
g = torch.full((1000,), float(rank + 1)) / world
red = GradAllReducer(bucket_bytes=4 * 300)
for b, e in [(800, 1000), (500, 800), (200, 500), (0, 200)]:
    red.ready(g, b, e)
red.finish()
expect = sum(r + 1 for r in range(world)) / world
assert torch.allclose(g, torch.full((1000,), expect)), g[:4]
assert 2 <= red.n_collectives <= 4
params -= 0.1 * g
assert replicas_consistent([params])
if rank == 1:
    params[3] += 1e-3
assert not replicas_consistent([params])
# callback_for(model): trunk stages bucketed, the last trunk stage flushes the rest (so the trunk travels during the towers' backward),
# the encoder's FC range and conv range go out as two immediate collectives
g1, g2 = torch.full((1000,), float(rank + 1)) / world, torch.full((600,), float(2 * rank + 1)) / world
class _DT: n_stages = 3; flat_grads = g1
class _Enc:
    flat_grads = g2
    def wait_fc_grads(self, stream): raise AssertionError("CPU tensors: no stream to wait on")
class _Model: dt = _DT(); encoder = _Enc()
red2 = GradAllReducer(bucket_bytes=4 * 5000)
cb = red2.callback_for(_Model())
cb(0, 700, 1000); cb(1, 300, 700)
assert red2.n_collectives == 0 and red2._pending
cb(2, 0, 300)
assert red2.n_collectives == 1 and not red2._pending
cb("encoder_fc", 250, 600); cb("encoder", 0, 250)
assert red2.n_collectives == 3
red2.finish()
assert torch.allclose(g1, torch.full((1000,), expect)) and torch.allclose(g2, torch.full((600,), sum(2 * r + 1 for r in range(world)) / world))
dist.destroy_process_group()
print("ok", rank)
"""


def test_data_parallel_allreduce_world2_gloo(tmp_path):
    script = tmp_path / "dp_check.py"
    script.write_text(_DP_SCRIPT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29731", str(script), ROOT], capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


def test_no_global_access_is_hoisted_above_the_pdl_wait():
    """Every kernel launched with programmatic dependent launch starts while its predecessor may still be running and must not touch
    global memory before `griddepcontrol.wait` (SASS: ACQBULK).  nvcc is free to hoist `const __restrict__` / __ldg loads above the
    asm statement (it did, in opt_adamw_kernel: LDG.E.CONSTANT of the step scalars before ACQBULK -> stale clip coefficient), so the
    compiled SASS of every kernel that contains the wait is scanned: only shared-memory traffic may precede it."""
    import glob
    import re
    import shutil
    import subprocess

    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    objs = sorted(glob.glob(os.path.join(ROOT, "multimodal-transformer-rl_b200", "build", "*.o")))
    if not objs:
        pytest.skip("library objects not built (run __graft_entry__.build())")
    offenders, checked = [], 0
    for obj in objs:
        sass = subprocess.run([cuobjdump, "-sass", obj], capture_output=True, text=True, check=True).stdout
        for body in re.split(r"Function : ", sass)[1:]:
            name = body.split()[0]
            if "ACQBULK" not in body:
                continue  # launched without the PDL attribute (plain stream order)
            checked += 1
            before = body.split("ACQBULK")[0]
            ops = re.findall(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", before)
            bad = [o for o in ops if o.startswith(("LDG", "STG", "ATOMG", "RED", "LDTM", "UTMALDG", "UBLKCP"))]
            if bad:
                offenders.append((os.path.basename(obj), name[:60], bad))
    assert checked >= 20, checked
    assert not offenders, offenders


def test_modal_attention_host_layout_and_no_fallback():
    """attn.* parameters carry the reference's names / shapes (agents.py:61-64, 81-82), a reference-style state_dict round-trips, the
    C layout functions agree with them, and nothing computes without the GPU."""
    import ctypes as C

    from mtrl_b200 import _lib
    from mtrl_b200.encoder import MultimodalStateEncoder
    from oracle.encoders import MultimodalStateEncoder as OracleEncoder

    lib = _lib.load()
    for attn, conv, fusion, sd in [("casl", "big", "concat", 1024), ("new", "small", "sum", 256)]:
        e = MultimodalStateEncoder((1, 84, 84), (1, 84, 84), conv, fusion, device="cpu", attn_type=attn)
        o = OracleEncoder((1, 84, 84), (1, 84, 84), conv, fusion, attn_type=attn)
        assert e.state_dim == sd
        assert {k: tuple(v.shape) for k, v in e.state_dict().items()} == {k: tuple(v.shape) for k, v in o.state_dict().items()}
        e.load_state_dict(o.state_dict())
        for k, v in o.state_dict().items():
            assert torch.equal(e.state_dict()[k], v), k
        # every parameter is a view into the one flat buffer the optimizer / all-reduce see
        base = e.flat_params.data_ptr()
        for _, p in e.named_parameters():
            assert base <= p.data_ptr() < base + e.flat_params.numel() * 4
        with pytest.raises(_lib.DtbError):
            e(torch.zeros(1, 1, 1, 84, 84), torch.zeros(1, 1, 1, 84, 84))
    with pytest.raises(NotImplementedError):
        MultimodalStateEncoder((1, 84, 84), (1, 84, 84), device="cpu", attn_type="other")  # as MinecraftAgent.__init__ (agents.py:114)
    bad = _lib.ModalAttnConfig(2, 512, 0)
    assert lib.dtb_modal_attn_param_entries(C.byref(bad)) < 0 and b"type" in lib.dtb_last_error()
    bad = _lib.ModalAttnConfig(0, 100, 0)
    assert lib.dtb_modal_attn_workspace_bytes(C.byref(bad), 8) == 0


def test_mfcc_host_geometry_and_no_fallback():
    """Frame count / FFT size / filter-bank bins come from host code in the library (no GPU needed) and must equal the oracle's for the
    reference's two asset sample rates; compute entry points refuse to run without the device."""
    import ctypes as C

    from mtrl_b200 import _lib, audio
    from oracle import mfcc as O

    lib = _lib.load()
    for sr, n in [(44100, 768764), (24000, 2 * 118203)]:  # noise.wav; target_good.wav (stereo, interleaved)
        assert lib.dtb_mfcc_num_frames(n, sr) == O.frame_geometry(n, sr)[2]
        nfft, bins = C.c_int(), (C.c_int * 28)()
        _lib.check(lib.dtb_mfcc_filterbank_bins(sr, C.byref(nfft), bins))
        assert nfft.value == O.calculate_nfft(sr, 0.025)
        assert list(bins) == [int(b) for b in O.filterbank_bins(26, nfft.value, sr)]
        assert lib.dtb_mfcc_workspace_bytes(1, n, sr, 84) > 13 * 8 * O.frame_geometry(n, sr)[2]
    assert lib.dtb_mfcc_num_frames(1000, 500) < 0  # unsupported sample rate: error code, not a crash
    assert lib.dtb_mfcc_num_frames(1000, 400000) < 0  # FFT would exceed the shared-memory kernel's 4096 points
    with pytest.raises(_lib.DtbError):
        audio.wav_to_mfcc(16000, torch.zeros(16000, dtype=torch.int16), device="cpu")
    with pytest.raises(_lib.DtbError):
        audio.resize_normalize(torch.zeros(50, 13, dtype=torch.float64))


def test_parameter_updates_after_to_are_seen_by_the_staleness_stamp():
    """``model.to(dev)`` rebinds every Parameter with ``param.data = view`` (own version counters), so an in-place
    update through the Parameters (torch.optim.AdamW) no longer bumps ``_flat._version``; the bf16 re-staging must still notice it."""
    from mtrl_b200.encoder import MultimodalStateEncoder
    from mtrl_b200.model import DecisionTransformer

    dt = DecisionTransformer(11, 3, 64, max_length=5, max_ep_len=50, n_layer=1, n_head=1, device="cpu")
    enc = MultimodalStateEncoder((1, 36, 36), (1, 36, 36), "small", "concat", device="cpu")
    for mod in (dt, enc):
        mod = mod.to("cpu")
        before_flat, before = mod._flat.clone(), mod._version_now()
        opt = torch.optim.AdamW(mod.parameters(), lr=1e-2)
        for p in mod.parameters():
            p.grad = torch.ones_like(p)
        opt.step()
        assert not torch.equal(mod._flat, before_flat)  # the fp32 master did change (the views alias the new flat buffer)
        assert mod._version_now() != before              # ... and the stamp that drives the bf16 re-staging sees it
        # every Parameter is still a view of the flat buffer
        assert all(p.data_ptr() == mod._flat.data_ptr() + 4 * off for p, off in zip(mod._plist, mod._offsets))


def test_fused_optimizer_zero_grad_attaches_and_clears_flat_views():
    """zero_grad() must leave every param.grad as a zeroed view of flat_grads (autograd then accumulates where step() reads);
    a grad living elsewhere is refused by step()."""
    from mtrl_b200.model import DecisionTransformer
    from mtrl_b200.optim import FusedClipAdamW

    dt = DecisionTransformer(11, 3, 64, max_length=5, max_ep_len=50, n_layer=1, n_head=1, device="cpu")
    opt = FusedClipAdamW([dt])
    dt.flat_grads.fill_(3.0)
    opt.zero_grad()
    assert float(dt.flat_grads.abs().sum()) == 0.0
    for p, off in zip(dt._plist, dt._offsets):
        assert p.grad is not None and p.grad.data_ptr() == dt.flat_grads.data_ptr() + 4 * off
    opt._check_grads_alias()
    dt._plist[3].grad = torch.zeros_like(dt._plist[3])  # what autograd does when .grad was None
    with pytest.raises(RuntimeError):
        opt._check_grads_alias()


def test_reset_parameters_follows_upstream_init_split():
    """GPT-2 init only inside ``transformer.*``; embeddings / heads keep torch's nn.Linear / nn.Embedding defaults."""
    from mtrl_b200.model import DecisionTransformer

    torch.manual_seed(0)
    dt = DecisionTransformer(11, 3, 128, max_length=5, max_ep_len=500, n_layer=2, n_head=1, device="cpu")
    sd = dt.state_dict()
    assert abs(float(sd["embed_timestep.weight"].std()) - 1.0) < 0.05                      # nn.Embedding: N(0, 1)
    bound = 1.0 / 11 ** 0.5
    assert float(sd["embed_state.weight"].abs().max()) <= bound + 1e-6                     # kaiming_uniform(a = sqrt 5): U(+-1/sqrt(fan_in))
    assert float(sd["embed_state.bias"].abs().max()) <= bound + 1e-6 and float(sd["embed_state.bias"].abs().sum()) > 0
    assert abs(float(sd["transformer.h.0.attn.c_attn.weight"].std()) - 0.02) < 2e-3        # GPT-2: N(0, 0.02)
    assert abs(float(sd["transformer.h.1.mlp.c_proj.weight"].std()) - 0.02 / 2.0) < 2e-3   # scaled by 1 / sqrt(2 L)
    assert float(sd["transformer.h.0.mlp.c_fc.bias"].abs().sum()) == 0.0
    assert torch.equal(sd["embed_ln.weight"], torch.ones(128)) and torch.equal(sd["transformer.ln_f.weight"], torch.ones(128))


def test_argument_validation_of_the_round2_entry_points_needs_no_gpu():
    """dtb_colsum_bf16_static / dtb_dt_decode_step / dtb_encoder_wait_fc_grads reject bad arguments before any CUDA call."""
    import ctypes as C

    from mtrl_b200 import _lib
    from mtrl_b200.model import DecisionTransformer

    lib = _lib.load()
    fake_x, fake_out = C.c_void_p(0x10000), C.c_void_p(0x20000)  # never dereferenced: the calls below are rejected first
    assert lib.dtb_colsum_bf16_static(None, fake_x, 16, 1032, 1032, fake_out, 0) == -1
    assert b"exceeds the static scratch limit" in lib.dtb_last_error()
    assert lib.dtb_colsum_bf16_static(None, fake_x, 16, 12, 16, fake_out, 0) == -1      # N not a multiple of 8
    assert lib.dtb_colsum_bf16_static(None, None, 16, 8, 8, fake_out, 0) == -1           # null input
    dt = DecisionTransformer(11, 3, 128, max_length=20, max_ep_len=100, device="cpu")
    assert lib.dtb_dt_decode_supported(dt._h) == 2                                          # hidden 128: the tensor-core decode kernel
    assert lib.dtb_dt_decode_step(dt._h, None, None, None, None, None, None, None, None, None, 20, None, 0, None) == -1
    wide = DecisionTransformer(11, 3, 64, max_length=20, max_ep_len=100, n_head=2, device="cpu")
    assert lib.dtb_dt_decode_supported(wide._h) == 1                                        # other widths: the CUDA-core cluster kernel
    assert dt._use_cache(None) is False      # training mode (dropout 0.1 active): never cached
    dt.eval(); wide.eval()
    assert dt._use_cache(None) is True and wide._use_cache(None) is False and wide._use_cache(True) is True  # automatic choice = tensor-core kernel only


def test_traffic_table_is_reproducible_from_the_committed_csv(tmp_path):
    """bench.py's roofline.traffic comes from profiles/r02_traffic.json; that file must be exactly what tools/ncu_traffic.py makes of the
    committed ncu CSV ('recomputable by running one committed script on one committed CSV')."""
    import importlib.util
    import json
    import shutil

    committed = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
    work = tmp_path / "repo"
    (work / "profiles").mkdir(parents=True)
    (work / "tools").mkdir()
    shutil.copy(os.path.join(ROOT, "tools", "ncu_traffic.py"), work / "tools" / "ncu_traffic.py")
    shutil.copy(os.path.join(ROOT, "profiles", "r02_c2_traffic_v3.csv"), work / "profiles" / "r02_c2_traffic_v3.csv")
    r = subprocess.run([sys.executable, str(work / "tools" / "ncu_traffic.py"), "c2", str(work / "profiles" / "r02_c2_traffic_v3.csv")],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    again = json.load(open(work / "profiles" / "r02_traffic.json"))
    assert set(again["c2"]) == set(committed["c2"])
    for fam, row in committed["c2"].items():
        for k in ("dram_bytes_per_launch", "dram_read_per_launch", "dram_write_per_launch", "launches_in_capture"):
            assert again["c2"][fam][k] == pytest.approx(row[k], rel=1e-12), (fam, k)
    gemm = committed["c2"]["gemm_tcgen05_kernel, HBM-bound shapes"]
    assert 4e6 < gemm["dram_bytes_per_launch"] < 6e6  # 4.97 MB per launch: below the 7.70 MB algorithmic bytes (no re-reads)


def test_bench_minecraft_audio_is_drawn_from_a_three_image_table():
    """bench.make_batch mirrors the reference environment: every timestep's spectrogram of a mono-audio workload is one of three images
    (REF environments/Minecraft/Minecraft.py:145-159); the index form expands to exactly the full tensor, stereo workloads stay free-form."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import bench

    wl = dict(bench.WORKLOADS["c2"])
    a, b = bench.make_batch(wl, 4, seed=5), bench.make_batch(wl, 4, seed=5, with_index=True)
    assert set(b) - set(a) == {"audio_index", "audio_table"} and all(torch.equal(a[k], b[k]) for k in a)
    assert b["audio_table"].shape == (3, 1, 64, 64) and b["audio_index"].dtype == torch.long and int(b["audio_index"].max()) <= 2
    assert torch.equal(b["audio_table"][b["audio_index"]], b["audio"])
    assert float(b["audio"].min()) >= -1.0 and float(b["audio"].max()) <= 1.0
    assert "audio_index" not in bench.make_batch(dict(bench.WORKLOADS["c3"]), 2, seed=5, with_index=True)
