"""ctypes binding of libdtb200.so (the C ABI in include/dtb200.h).  Fails loudly when the library is absent."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdtb200.so")

_lib = None


class DtbError(RuntimeError):
    pass


class GemmEpilogue(C.Structure):
    _fields_ = [
        ("bias", C.c_void_p), ("residual", C.c_void_p), ("ldr", C.c_int),
        ("aux", C.c_void_p), ("ldaux", C.c_int), ("act", C.c_int), ("dact", C.c_int),
        ("out_bf16", C.c_void_p), ("ldo_b", C.c_int), ("out_f32", C.c_void_p), ("ldo_f", C.c_int),
        ("out_preact", C.c_void_p), ("partial", C.c_void_p), ("alpha", C.c_float), ("accumulate", C.c_int),
    ]


class DtConfig(C.Structure):
    _fields_ = [("state_dim", C.c_int), ("act_dim", C.c_int), ("hidden", C.c_int), ("n_layer", C.c_int), ("n_head", C.c_int),
                ("n_inner", C.c_int), ("max_ep_len", C.c_int), ("activation", C.c_int), ("action_tanh", C.c_int), ("ln_eps", C.c_float)]


class DtBatch(C.Structure):
    _fields_ = [("B", C.c_int), ("K", C.c_int), ("states", C.c_void_p), ("states_bf16", C.c_void_p), ("actions", C.c_void_p),
                ("returns_to_go", C.c_void_p), ("timesteps", C.c_void_p), ("attention_mask", C.c_void_p), ("states_relu", C.c_int)]


class AdamwSegment(C.Structure):
    _fields_ = [("params", C.c_void_p), ("grads", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("params_bf16", C.c_void_p), ("n", C.c_longlong)]


class AdamwHparams(C.Structure):
    _fields_ = [("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double), ("weight_decay", C.c_double),
                ("max_norm", C.c_double), ("lr_dev", C.c_void_p), ("write_clipped_grads", C.c_int), ("warmup_steps", C.c_int)]


class TrajStore(C.Structure):
    _fields_ = [("obs", C.c_void_p), ("state_dim", C.c_int), ("state_mean", C.c_void_p), ("state_std", C.c_void_p), ("act", C.c_void_p),
                ("act_dim", C.c_int), ("rew", C.c_void_p), ("done", C.c_void_p), ("rtg_suffix", C.c_void_p), ("traj_offset", C.c_void_p),
                ("n_raw", C.c_int), ("raw_src", C.c_void_p * 4), ("raw_row_bytes", C.c_int * 4)]


class BatchOut(C.Structure):
    _fields_ = [("states", C.c_void_p), ("actions", C.c_void_p), ("rewards", C.c_void_p), ("dones", C.c_void_p), ("rtg", C.c_void_p),
                ("timesteps", C.c_void_p), ("mask", C.c_void_p), ("raw_dst", C.c_void_p * 4)]


class TowerConfig(C.Structure):
    _fields_ = [("in_channels", C.c_int), ("in_h", C.c_int), ("in_w", C.c_int), ("big", C.c_int)]


class EncoderConfig(C.Structure):
    _fields_ = [("video", TowerConfig), ("audio", TowerConfig), ("fusion", C.c_int), ("input_scale", C.c_float)]


class ModalAttnConfig(C.Structure):
    _fields_ = [("type", C.c_int), ("feat", C.c_int), ("fusion", C.c_int)]


def _declare(lib):
    """restype / argtypes for the entry points whose defaults (int) would truncate pointers or sizes."""
    vp, i, f, sz, ll = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_longlong
    lib.dtb_last_error.restype = C.c_char_p
    lib.dtb_abi_version.restype = i
    lib.dtb_launch_count.restype = ll
    lib.dtb_set_option.argtypes = [C.c_char_p, i]
    lib.dtb_profile_set_balance.argtypes = [C.c_double]
    lib.dtb_profile_set_balance.restype = None
    lib.dtb_profile_enable.argtypes = [i]
    lib.dtb_profile_enable.restype = None
    lib.dtb_profile_span.argtypes = [i, C.POINTER(i), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.dtb_profile_report.argtypes = [i, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(ll)]
    for name in ("dtb_layernorm_bwd_scratch_floats", "dtb_colsum_scratch_floats", "dtb_mse_scratch_floats"):
        getattr(lib, name).restype = sz
    lib.dtb_dt_create.argtypes = [C.POINTER(DtConfig), C.POINTER(vp)]
    lib.dtb_dt_destroy.argtypes = [vp]
    lib.dtb_dt_param_entries.argtypes = [vp]
    lib.dtb_dt_param_layout.argtypes = [vp, C.POINTER(ll), C.POINTER(ll)]
    lib.dtb_dt_param_layout.restype = ll
    lib.dtb_dt_workspace_bytes.argtypes = [vp, i, i]
    lib.dtb_dt_workspace_bytes.restype = sz
    lib.dtb_dt_stage_params.argtypes = [vp, vp, vp, vp]
    lib.dtb_dt_forward.argtypes = [vp, vp, vp, vp, C.POINTER(DtBatch), vp, sz, vp, vp, vp]
    lib.dtb_dt_loss.argtypes = [vp, vp, C.POINTER(DtBatch), vp, f, vp, sz, vp]
    lib.dtb_dt_backward.argtypes = [vp, vp, vp, vp, C.POINTER(DtBatch), vp, sz, vp, vp, vp, vp, vp, i, i]
    lib.dtb_dt_stage_grad_range.argtypes = [vp, i, C.POINTER(ll), C.POINTER(ll)]
    lib.dtb_dt_decode_supported.argtypes = [vp]
    lib.dtb_dt_decode_cache_bytes.argtypes = [vp, i]
    lib.dtb_dt_decode_cache_bytes.restype = sz
    lib.dtb_dt_decode_stage.argtypes = [vp, vp, vp, i, vp, sz]
    lib.dtb_dt_decode.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, vp, sz, vp]
    lib.dtb_dt_decode_step.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, vp, sz, vp]
    lib.dtb_encoder_create.argtypes = [C.POINTER(EncoderConfig), C.POINTER(vp)]
    lib.dtb_encoder_destroy.argtypes = [vp]
    lib.dtb_encoder_state_dim.argtypes = [vp]
    lib.dtb_encoder_param_entries.argtypes = [vp]
    lib.dtb_encoder_param_layout.argtypes = [vp, C.POINTER(ll), C.POINTER(ll)]
    lib.dtb_encoder_param_layout.restype = ll
    lib.dtb_encoder_workspace_bytes.argtypes = [vp, i]
    lib.dtb_encoder_workspace_bytes.restype = sz
    lib.dtb_encoder_stage_params.argtypes = [vp, vp, vp, vp]
    lib.dtb_encoder_staged_elems.argtypes = [vp]
    lib.dtb_encoder_fc_grad_range.argtypes = [vp, C.POINTER(ll), C.POINTER(ll)]
    lib.dtb_encoder_wait_fc_grads.argtypes = [vp, vp]
    lib.dtb_encoder_staged_elems.restype = ll
    lib.dtb_encoder_forward.argtypes = [vp, vp, vp, vp, i, vp, vp, vp, sz, vp]
    lib.dtb_encoder_backward.argtypes = [vp, vp, vp, vp, i, vp, vp, vp, vp, i, vp, sz, vp]
    lib.dtb_encoder_forward_u8.argtypes = lib.dtb_encoder_forward.argtypes
    lib.dtb_encoder_backward_u8.argtypes = lib.dtb_encoder_backward.argtypes
    macp = C.POINTER(ModalAttnConfig)
    lib.dtb_modal_attn_param_entries.argtypes = [macp]
    lib.dtb_modal_attn_param_layout.argtypes = [macp, C.POINTER(ll), C.POINTER(ll)]
    lib.dtb_modal_attn_param_layout.restype = ll
    lib.dtb_modal_attn_state_dim.argtypes = [macp]
    lib.dtb_modal_attn_workspace_bytes.argtypes = [macp, i]
    lib.dtb_modal_attn_workspace_bytes.restype = sz
    lib.dtb_modal_attn_forward.argtypes = [macp, vp, vp, i, vp, vp, vp, sz, vp, vp]
    lib.dtb_modal_attn_backward.argtypes = [macp, vp, vp, i, vp, i, vp, vp, sz, vp, vp]
    lib.dtb_mfcc_num_frames.argtypes = [ll, i]
    lib.dtb_mfcc_num_frames.restype = ll
    lib.dtb_mfcc_filterbank_bins.argtypes = [i, C.POINTER(i), C.POINTER(i)]
    lib.dtb_mfcc_workspace_bytes.argtypes = [i, ll, i, i]
    lib.dtb_mfcc_workspace_bytes.restype = sz
    lib.dtb_mfcc_cepstra.argtypes = [vp, vp, i, i, i, ll, ll, i, vp, vp]
    lib.dtb_mfcc_table_doubles.argtypes = [i]
    lib.dtb_mfcc_table_doubles.restype = sz
    lib.dtb_mfcc_resize_normalize.argtypes = [vp, vp, i, ll, i, i, vp, vp]
    lib.dtb_wav_to_mfcc_image.argtypes = [vp, vp, i, i, i, ll, ll, i, i, i, vp, sz, vp, vp]
    lib.dtb_mfcc_build_tables.argtypes = [vp, i, vp]
    lib.dtb_pad_left.argtypes = [vp, i, C.POINTER(vp), C.POINTER(vp), C.POINTER(i), i, i, vp]
    lib.dtb_dt_set_dropout.argtypes = [vp, f, f, f, vp]
    lib.dtb_get_batch.argtypes = [vp, C.POINTER(TrajStore), vp, vp, i, i, i, C.c_double, C.POINTER(BatchOut)]
    lib.dtb_adamw_scratch_floats.restype = sz
    lib.dtb_clip_adamw_step.argtypes = [vp, i, C.POINTER(AdamwSegment), C.POINTER(AdamwHparams), vp, vp]
    lib.dtb_attention_fwd_dropout.argtypes = [vp, vp, vp, i, i, i, i, vp, vp, vp, C.c_uint, f]
    lib.dtb_attention_bwd_dropout.argtypes = [vp, vp, vp, vp, vp, vp, i, i, i, i, vp, vp, vp, C.c_uint, f]
    lib.dtb_dropout_advance.argtypes = [vp, vp]
    lib.dtb_dropout_mask.argtypes = [vp, vp, C.c_uint, f, i, ll, vp]
    lib.dtb_masked_mse.argtypes = [vp, vp, vp, vp, i, i, f, vp, vp, vp]
    lib.dtb_heads_fwd.argtypes = [vp] * 8 + [i] * 5 + [vp] * 3
    lib.dtb_embed_fwd.argtypes = [vp] * 15 + [i] * 5 + [f] + [vp] * 5 + [i, i]


def load() -> C.CDLL:
    """Load libdtb200.so; raise if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DtbError(
            f"{LIB_PATH} not found: build it with `python multimodal-transformer-rl_b200/build.py` "
            "(the sm_100a CUDA extension is the only implementation; there is no CPU/PyTorch fallback)"
        )
    lib = C.CDLL(LIB_PATH)
    _declare(lib)
    _lib = lib
    return lib


def check(rc: int) -> int:
    if rc < 0:
        raise DtbError(f"dtb200 error {rc}: {load().dtb_last_error().decode()}")
    return rc


def ptr(t) -> C.c_void_p:
    """Device pointer of a torch tensor (None -> NULL)."""
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr() -> C.c_void_p:
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def exported_symbols() -> list[str]:
    """Names declared in include/dtb200.h (parsed) — used by the CPU test that the ABI is complete."""
    import re

    hdr = os.path.join(os.path.dirname(_HERE), "include", "dtb200.h")
    txt = open(hdr).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(dtb_[a-z0-9_]+)\s*\(", txt)))


def gemm_bf16(A, B, M, N, K, *, a_mn_major=False, b_mn_major=False, bias=None, residual=None, aux=None, act=0, dact=0,
              out_bf16=None, out_f32=None, out_preact=None, partial=None, alpha=1.0, accumulate=False, split_k=1, bn=0):
    """Raw tensor-core GEMM through the C ABI (see dtb_gemm_bf16 in include/dtb200.h)."""
    ep = GemmEpilogue()
    ep.bias = ptr(bias).value
    ep.residual = ptr(residual).value
    ep.ldr = residual.stride(0) if residual is not None else 0
    ep.aux = ptr(aux).value
    ep.ldaux = aux.stride(0) if aux is not None else 0
    ep.act, ep.dact = act, dact
    ep.out_bf16 = ptr(out_bf16).value
    ep.ldo_b = out_bf16.stride(0) if out_bf16 is not None else (out_preact.stride(0) if out_preact is not None else 0)
    ep.out_f32 = ptr(out_f32).value
    ep.ldo_f = out_f32.stride(0) if out_f32 is not None else 0
    ep.out_preact = ptr(out_preact).value
    ep.partial = ptr(partial).value
    ep.alpha = alpha
    ep.accumulate = int(accumulate)
    check(load().dtb_gemm_bf16(stream_ptr(), ptr(A), C.c_int(A.stride(0)), int(a_mn_major), ptr(B), C.c_int(B.stride(0)),
                               int(b_mn_major), M, N, K, C.byref(ep), split_k, bn))


def gemm_bf16_ln(A, B, M, N, K, gamma, beta, eps, ln_out, *, b_mn_major=False, bias=None, residual=None, out_f32=None, mean=None, rstd=None):
    """GEMM + bias + residual + fused LayerNorm epilogue (dtb_gemm_bf16_ln)."""
    ep = GemmEpilogue()
    ep.bias = ptr(bias).value
    ep.residual = ptr(residual).value
    ep.ldr = residual.stride(0) if residual is not None else 0
    ep.out_f32 = ptr(out_f32).value
    ep.ldo_f = out_f32.stride(0) if out_f32 is not None else 0
    ep.alpha = 1.0
    check(load().dtb_gemm_bf16_ln(stream_ptr(), ptr(A), C.c_int(A.stride(0)), 0, ptr(B), C.c_int(B.stride(0)), int(b_mn_major), M, N, K,
                                  C.byref(ep), ptr(gamma), ptr(beta), C.c_float(eps), ptr(ln_out), C.c_int(ln_out.stride(0)), ptr(mean),
                                  ptr(rstd)))


def set_option(name: str, value: int):
    """Runtime switch between kernel forms (see dtb_set_option in include/dtb200.h)."""
    check(load().dtb_set_option(name.encode(), int(value)))


# ---- thin op wrappers (used by tests and by the host-side model) ------------------------------------------------------
def _f32p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def layernorm_fwd(x, gamma, beta, eps, y, mean=None, rstd=None):
    M, H = x.shape
    check(load().dtb_layernorm_fwd(stream_ptr(), ptr(x), ptr(gamma), ptr(beta), M, H, C.c_float(eps), ptr(y), ptr(mean), ptr(rstd)))


def layernorm_bwd(dy, x, mean, rstd, gamma, dres, dx_f32, dx_bf16, dgamma, dbeta, scratch=None, dcolsum=None):
    import torch

    M, H = x.shape
    lib = load()
    lib.dtb_layernorm_bwd_scratch_floats.restype = C.c_size_t
    if scratch is None:
        scratch = torch.empty(lib.dtb_layernorm_bwd_scratch_floats(H), device=x.device, dtype=torch.float32)
    if dcolsum is not None:
        check(lib.dtb_layernorm_bwd_colsum(stream_ptr(), ptr(dy), ptr(x), ptr(mean), ptr(rstd), ptr(gamma), ptr(dres), M, H, ptr(dx_f32),
                                           ptr(dx_bf16), ptr(dgamma), ptr(dbeta), ptr(dcolsum), ptr(scratch)))
        return
    check(lib.dtb_layernorm_bwd(stream_ptr(), ptr(dy), ptr(x), ptr(mean), ptr(rstd), ptr(gamma), ptr(dres), M, H, ptr(dx_f32),
                                ptr(dx_bf16), ptr(dgamma), ptr(dbeta), ptr(scratch)))


def attention_fwd(qkv, mask, n_head, out, lse, drop_state=None, site=0, p=0.0):
    B, T, H3 = qkv.shape
    d = H3 // 3 // n_head
    if drop_state is not None and p > 0:
        check(load().dtb_attention_fwd_dropout(stream_ptr(), ptr(qkv), ptr(mask), B, T, n_head, d, ptr(out), ptr(lse), ptr(drop_state),
                                               int(site), float(p)))
        return
    check(load().dtb_attention_fwd(stream_ptr(), ptr(qkv), ptr(mask), B, T, n_head, d, ptr(out), ptr(lse)))


def attention_bwd(qkv, mask, out, dout, lse, n_head, dqkv, delta, drop_state=None, site=0, p=0.0):
    B, T, H3 = qkv.shape
    d = H3 // 3 // n_head
    if drop_state is not None and p > 0:
        check(load().dtb_attention_bwd_dropout(stream_ptr(), ptr(qkv), ptr(mask), ptr(out), ptr(dout), ptr(lse), B, T, n_head, d,
                                               ptr(dqkv), ptr(delta), ptr(drop_state), int(site), float(p)))
        return
    check(load().dtb_attention_bwd(stream_ptr(), ptr(qkv), ptr(mask), ptr(out), ptr(dout), ptr(lse), B, T, n_head, d, ptr(dqkv),
                                   ptr(delta)))


def cast_transpose(src, dst, dst_t):
    R, Cc = src.shape
    check(load().dtb_cast_transpose(stream_ptr(), ptr(src), R, Cc, ptr(dst), ptr(dst_t)))


def colsum_bf16(x, out, accumulate=False):
    import torch

    M, N = x.shape
    lib = load()
    lib.dtb_colsum_scratch_floats.restype = C.c_size_t
    scratch = torch.empty(lib.dtb_colsum_scratch_floats(N), device=x.device, dtype=torch.float32)
    check(lib.dtb_colsum_bf16(stream_ptr(), ptr(x), M, N, x.stride(0), ptr(out), int(accumulate), ptr(scratch)))


def colsum_bf16_static(x, out, accumulate=False):
    """dtb_colsum_bf16_static: same result as colsum_bf16, scratch in the library's static device array; allocates nothing (N <= 1024)."""
    M, N = x.shape
    check(load().dtb_colsum_bf16_static(stream_ptr(), ptr(x), M, N, x.stride(0), ptr(out), int(accumulate)))


def embed_fwd(states, state_proj, actions, rtg, timesteps, table, wR, bR, Ws, bs, Wa, ba, gamma, beta, eps, x, pre=None, mean=None,
              rstd=None, U=None, u_states=True):
    BK, ad = actions.shape
    H = table.shape[1]
    sd = Ws.shape[1]
    P8 = U.shape[1] if U is not None else 0
    check(load().dtb_embed_fwd(stream_ptr(), ptr(states), ptr(state_proj), ptr(actions), ptr(rtg), ptr(timesteps), ptr(table), ptr(wR),
                               ptr(bR), ptr(Ws), ptr(bs), ptr(Wa), ptr(ba), ptr(gamma), ptr(beta), BK, sd, ad, H, table.shape[0],
                               float(eps), ptr(x), ptr(pre), ptr(mean), ptr(rstd), ptr(U), P8, int(u_states)))


def heads_fwd(y, Wa, ba, Ws, bs, Wr, br, action_tanh, action_preds, state_preds=None, return_preds=None):
    BK, ad = action_preds.shape
    H = y.shape[1]
    sd = Ws.shape[0] if Ws is not None else 0
    check(load().dtb_heads_fwd(stream_ptr(), ptr(y), ptr(Wa), ptr(ba), ptr(Ws), ptr(bs), ptr(Wr), ptr(br), BK, H, sd, ad, int(action_tanh),
                               ptr(action_preds), ptr(state_preds), ptr(return_preds)))


def masked_mse(pred, target, mask, grad_scale, loss_out, d_pred=None):
    import torch

    BK, ad = pred.shape
    lib = load()
    scratch = torch.empty(lib.dtb_mse_scratch_floats(), device=pred.device, dtype=torch.float32)
    check(lib.dtb_masked_mse(stream_ptr(), ptr(pred), ptr(target), ptr(mask), BK, ad, float(grad_scale), ptr(loss_out), ptr(d_pred),
                             ptr(scratch)))
