"""ctypes binding of libapplestar_b200.so (the C-ABI declared in include/applestar_b200.h).

The library is built in-tree by `__graft_entry__.build()` / `make -C applestar_b200/csrc`.  There is no
fallback: a missing library raises, a non-zero return code raises.
"""
import ctypes
import os
from ctypes import c_float, c_int, c_int32, c_int64, c_void_p, POINTER

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libapplestar_b200.so')
_lib = None

AS_LOSS_MAX_FIELDS = 6
AS_LOSS_MAX_STATS = 96

_ERR = {-1: 'AS_EINVAL (bad pointer / argument)', -2: 'AS_ESIZE (size out of range)',
        -3: 'AS_EALIGN (pointer or pitch not 16-byte aligned)', -4: 'AS_EUNSUPPORTED', -5: 'AS_EDRIVER (tensor-map encode failed)'}


class AsSuSampleArgs(ctypes.Structure):
    """as_su_sample_args (include/applestar_b200.h)."""
    PTRS = ('key', 'c0', 'c1', 'mt', 'wf1t', 'bf1', 'wq2t', 'bq2', 'wih_t', 'whh_t', 'ln_i_w', 'ln_i_b', 'ln_h_w', 'ln_h_b',
            'ln_c_w', 'ln_c_b', 'entity_num', 'su_mask', 'noise', 'logits', 'units', 'num', 'extra', 'emb', 'steps_row', 'ended_rows')
    _fields_ = [(n, c_void_p) for n in PTRS] + [('R', c_int32), ('N1', c_int32), ('S', c_int32), ('extra_units', c_int32),
                                                  ('temperature', c_float)]


class AsLossArgs(ctypes.Structure):
    _fields_ = [
        ('T', c_int32), ('B', c_int32), ('F', c_int32), ('S', c_int32), ('N1', c_int32), ('NE', c_int32),
        ('head_k', c_int32 * 6), ('upgo_field', c_int32), ('at_kl_steps', c_float), ('dapo_steps', c_float),
        ('logit', c_void_p * 6), ('teacher', c_void_p * 6), ('successive', c_void_p * 6), ('dlogit', c_void_p * 6), ('action', c_void_p * 6),
        ('mu_logp', c_void_p * 6), ('actions_mask', c_void_p * 6),
        ('su_mask', c_void_p), ('su_logits_mask', c_void_p), ('tu_logits_mask', c_void_p),
        ('cum_action_mask', c_void_p), ('step', c_void_p),
        ('value', c_void_p * AS_LOSS_MAX_FIELDS), ('dvalue', c_void_p * AS_LOSS_MAX_FIELDS),
        ('reward', c_void_p * AS_LOSS_MAX_FIELDS), ('field_mask', c_void_p * AS_LOSS_MAX_FIELDS),
        ('winloss_reward', c_void_p), ('battle_reward', c_void_p),
        ('gamma_td', c_float * AS_LOSS_MAX_FIELDS), ('w_pg', c_float * AS_LOSS_MAX_FIELDS),
        ('w_td', c_float * AS_LOSS_MAX_FIELDS),
        ('hw_pg', c_float * 6), ('hw_upgo', c_float * 6), ('hw_ent', c_float * 6), ('hw_kl', c_float * 6),
        ('hw_dapo', c_float * 6),
        ('w_upgo', c_float), ('w_ent', c_float), ('w_kl', c_float), ('w_atkl', c_float), ('w_dapo', c_float),
        ('stat_weight', c_void_p), ('grad_scale', c_void_p),
        ('row_stats', c_void_p * 6), ('coef', c_void_p * 6), ('su_step', c_void_p), ('partial', c_void_p), ('out', c_void_p),
    ]


_SIGS = {
    'as_abi_version': (c_int, []),
    'as_scatter_connection_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int,
                                          c_int, c_int, c_int, c_int, c_void_p]),
    'as_scatter_connection_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int,
                                          c_int, c_int, c_int, c_void_p]),
    'as_gemm_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                             c_int64, c_int64, c_int64, c_int, c_int, c_int64, c_int64, c_int64, c_int64, c_int64,
                             c_int64, c_int, c_int, c_float, c_int, c_void_p, c_int, c_void_p]),
    'as_gemm_bf16_relu_bits': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64, c_int64, c_void_p]),
    'as_gemm_bf16_bitmasked': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_int64, c_int64, c_int, c_float, c_void_p]),
    'as_gemm_bf16_masked': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_int64,
                                    c_int64, c_int, c_float, c_void_p]),
    'as_gemm_splits': (c_int, [c_int, c_int]),
    'as_splitk_reduce': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_conv_nhwc_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int,
                                  c_int, c_int, c_int, c_int, c_void_p]),
    'as_conv_wgrad_nhwc_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int,
                                        c_int, c_int, c_void_p]),
    'as_nchw_to_pnhwc': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    'as_pnhwc_to_nchw': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    'as_add_act_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_relu_bwd_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    'as_relu_bwd_f32': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    'as_gated_fwd_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    'as_gated_scratch_floats': (c_int64, []),
    'as_gated_bwd_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_int64, c_void_p]),
    'as_upsample2_fwd_bf16': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    'as_upsample2_bwd_bf16': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    'as_loc_cat_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    'as_loc_cat_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int,
                               c_void_p]),
    'as_conv_c32_1_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_float, c_void_p]),
    'as_conv_c32_1_scratch_floats': (c_int64, []),
    'as_conv_c32_1_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                  c_int, c_int, c_float, c_void_p]),
    'as_colsum_scratch_floats': (c_int64, [c_int]),
    'as_colsum_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_colsum_f32': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_colsum_bf16_static': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_su_prefix_fwd': (c_int, [c_void_p] * 7 + [c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    'as_su_prefix_bwd': (c_int, [c_void_p] * 6 + [c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    'as_su_mask_logits': (c_int, [c_void_p] * 4 + [c_int64, c_int, c_int, c_int, c_float, c_float, c_void_p]),
    'as_spatial_front_fwd': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_spatial_front_cells_bytes': (c_int64, [c_int64]),
    'as_spatial_front_bwd_scratch_floats': (c_int64, [c_int64, c_int]),
    'as_spatial_front_bwd': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                     c_int, c_void_p]),
    'as_upconv_c32_1_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_float, c_void_p]),
    'as_upconv_c32_1_scratch_floats': (c_int64, [c_int64, c_int]),
    'as_upconv_c32_1_bwd': (c_int, [c_void_p] * 7 + [c_int64, c_int, c_int, c_float, c_int, c_void_p]),
    'as_ln256_bf16_fwd': (c_int, [c_void_p] * 6 + [c_int64, c_int, c_void_p]),
    'as_ln256_bf16_bwd_scratch_floats': (c_int64, []),
    'as_ln256_bf16_bwd': (c_int, [c_void_p] * 9 + [c_int] + [c_void_p] * 4 + [c_int64, c_int, c_void_p]),
    'as_masked_mean256_bf16_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_masked_mean256_bf16_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_embed_relu_fwd': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_one_hot_bf16': (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_time_encode': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_bo_features': (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    'as_cast_any_bf16': (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_ce_rows_fwd': (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    'as_ce_rows_bwd': (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    'as_pointer_keys_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_pointer_keys_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_expand_rows_f32': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_void_p]),
    'as_conv_weight_layouts': (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    'as_spatial_fc_weight_fwd': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    'as_spatial_fc_weight_bwd': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    'as_pointer_logits_fwd': (c_int, [c_void_p] * 4 + [c_int64, c_int, c_int, c_int, c_float, c_float, c_void_p]),
    'as_pointer_logits_bwd': (c_int, [c_void_p] * 6 + [c_int64, c_int, c_int, c_int, c_float, c_void_p]),
    'as_attn_fwd_ok': (c_int, [c_int, c_int]),
    'as_attn_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_float, c_void_p]),
    'as_attn_bwd_dq': (c_int, [c_void_p] * 6 + [c_int64, c_int, c_int, c_int, c_float, c_void_p]),
    'as_attn_small_ok': (c_int, [c_int, c_int]),
    'as_attn_small_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_float, c_void_p]),
    'as_attn_small_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_float, c_void_p]),
    'as_pack_f32': (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    'as_conv_wgrad_pair_mode': (c_int, [c_int, c_int, c_int, c_int, c_int]),
    'as_conv_wgrad_ws_floats': (c_int64, [c_int, c_int, c_int, c_int, c_int]),
    'as_conv_nhwc_bf16_add': (c_int, [c_void_p] * 5 + [c_int64, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    'as_conv_nhwc_bf16_relu_bits': (c_int, [c_void_p] * 5 + [c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    'as_conv_nhwc_bf16_bitmasked': (c_int, [c_void_p] * 5 + [c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    'as_conv_nhwc_bf16_masked': (c_int, [c_void_p] * 5 + [c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    'as_su_sample': (c_int, [POINTER(AsSuSampleArgs), c_void_p]),
    'as_masked_mean_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_masked_mean_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_linear_bwd_prep': (c_int, [c_void_p] * 6 + [c_int64, c_int, c_void_p]),
    'as_linear_bwd_prep2': (c_int, [c_void_p] * 7 + [c_int64, c_int, c_void_p]),
    'as_relu_bwd_colsum_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_spatial_features': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                                    c_void_p]),
    'as_spatial_features_scatter': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int,
                                            c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_scatter_connection_bwd_pnhwc': (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64,
                                                c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    'as_maxpool2_fwd_bf16': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    'as_maxpool2_bwd_bf16': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    'as_layer_norm_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                  c_int, c_void_p]),
    'as_layer_norm_bwd_scratch_floats': (c_int64, [c_int]),
    'as_layer_norm_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_layer_norm_bwd_fused_scratch_floats': (c_int64, [c_int]),
    'as_layer_norm_bwd_fused': (c_int, [c_void_p] * 9 + [c_int] + [c_void_p] * 4 + [c_int64, c_int, c_int, c_void_p]),
    'as_lnlstm_fwd': (c_int, [c_void_p] * 13 + [c_int, c_int, c_int, c_void_p]),
    'as_lnlstm_bwd': (c_int, [c_void_p] * 21 + [c_int, c_int, c_int, c_void_p]),
    'as_lnlstm_cluster_ok': (c_int, [c_int]),
    'as_softmax_fwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    'as_softmax_bwd': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    'as_entity_features': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int64, c_int,
                                   c_void_p, c_void_p]),
    'as_cast_f32_bf16': (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    'as_gumbel_argmax': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    'as_rl_loss_forward': (c_int, [POINTER(AsLossArgs), c_void_p]),
    'as_rl_loss_backward': (c_int, [POINTER(AsLossArgs), c_float, c_void_p]),
    'as_debug_gemm_mode': (c_int, [c_int]),
    'as_debug_gemm_profile': (c_int, [c_void_p]),
    'as_debug_attn_profile': (c_int, [c_void_p]),
    'as_debug_lstm_profile': (c_int, [c_void_p]),
    'as_debug_lstm_max_clusters': (c_int, [c_int]),
    'as_clip_adam_scratch_bytes': (c_int64, []),
    'as_clip_adam_per_param': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float,
                                       c_float, c_float, c_float, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                       c_void_p]),
    'as_clip_adam': (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float,
                             c_float, c_float, c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
}

EXPORTS = tuple(_SIGS)
LAUNCHES = 0     # number of C-ABI kernel-launching calls made by this process (bench.py reports it)


def lib():
    """The loaded library; raises if it has not been built (no CPU / library fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError('libapplestar_b200.so not found at %s -- run `python -c "import __graft_entry__ as g; '
                               'g.build()"` (nvcc, sm_100a). There is no fallback path.' % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(rc, what):
    if rc != 0:
        msg = _ERR.get(rc, 'cudaError_t %d' % rc if rc > 0 else 'error %d' % rc)
        raise RuntimeError('%s failed: %s' % (what, msg))


def call(name, *args):
    """Call a kernel-launching entry point and raise on a non-zero return code."""
    global LAUNCHES
    LAUNCHES += 1
    check(getattr(lib(), name)(*args), name)


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream
