"""ctypes signatures of the libb200ps.so entry points (mirror of include/b200ps.h)."""
import ctypes as C

vp, i32, i64, u32, u64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_uint64, C.c_float

SIGNATURES = {
    "b200ps_mlp_tail": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, f32, i32, vp, i32, i64, vp],
    "b200ps_mlp_tail_static": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, f32, i32, vp, i32, i64, vp],
    "b200ps_softmax_ce": [vp, vp, vp, vp, vp, i32, i32, i64, i32, f32, f32, vp],
    "b200ps_colsum": [vp, vp, i32, i32, i64, i32, f32, vp],
    "b200ps_sgd_apply": [vp, vp, i32, vp, u64, f32, f32, vp],
    "b200ps_adam_apply": [vp, vp, vp, vp, u64, f32, f32, f32, f32, vp, vp],
    "b200ps_adam_advance": [vp, f32, f32, vp],
    "b200ps_linreg_fwd_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, vp],
    "b200ps_synth_mnist": [vp, vp, vp, i32, i32, i32, u64, u64, vp],
    "b200ps_synth_dense_bf16": [vp, vp, i32, i32, i32, u64, u64, vp],
    "b200ps_synth_sparse": [vp, vp, vp, vp, i32, i32, i32, i64, f32, u64, u64, vp],
    "b200ps_init_normal": [vp, u64, f32, f32, u64, u64, u64, vp],
    "b200ps_cast_f32_bf16": [vp, vp, u64, vp],
    "b200ps_mlp_fused_step_set_timing": [vp],
    "b200ps_mlp_fused_step": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32,
                              i32, f32, f32, f32, vp, vp],
    "b200ps_sparse_wd": [vp, vp, vp, vp, vp, i64, vp, i64, i32, vp, i64, vp, vp, vp, vp, vp, vp, i32, i32, i32,
                         f32, f32, f32, vp],
    "b200ps_sparse_lr": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i64, i32, i32, f32, f32, f32, vp],
    "b200ps_auc_hist": [vp, vp, i32, vp, i32, vp, vp, vp],
    "b200ps_push_sync_sgd": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i64, i32, i32, i32, f32, i32, vp, i32, i32, i32, vp],
    "b200ps_push_sync_ll": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i64, i64, f32, vp, i32, i32, vp],
    "b200ps_push_async": [vp, vp, vp, vp, i32, i32, i32, i64, f32, i32, vp, i32, i32, vp],
    "b200ps_pull": [vp, vp, vp, i32, i32, i32, i64, i32, i32, i32, vp],
    "b200ps_barrier": [vp, i32, i32, i32, vp],
    "b200ps_symm_alloc": [u64, C.POINTER(vp), vp],
    "b200ps_symm_open": [vp, C.POINTER(vp)],
    "b200ps_symm_close": [vp],
    "b200ps_symm_free": [vp],
    "b200ps_device_count": [C.POINTER(i32)],
    "b200ps_can_access_peer": [i32, i32, C.POINTER(i32)],
    "b200ps_host_step": [vp, vp, u64, vp, vp, u64, vp, vp, vp, u64, vp, vp],
    "b200ps_feeder_create": [C.POINTER(vp), i32, u64, u64, u64],
    "b200ps_feeder_result_ptr": [vp, i32, C.POINTER(vp)],
    "b200ps_feeder_bind": [vp, i32, vp, vp, vp],
    "b200ps_feeder_submit": [vp, vp, vp, vp, C.POINTER(i64)],
    "b200ps_feeder_wait": [vp, i64, vp],
    "b200ps_feeder_stats": [vp, C.POINTER(C.c_double), C.POINTER(i64)],
    "b200ps_feeder_destroy": [vp],
    "b200ps_event_create": [C.POINTER(vp)],
    "b200ps_event_synchronize": [vp],
    "b200ps_event_destroy": [vp],
    "b200ps_libsvm_count": [C.c_char_p, u64, C.POINTER(u64), C.POINTER(u64)],
}

# entry points whose return type is not the int status code
SPECIAL = {
    "b200ps_signal_pad_words": (C.c_int, []),
    "b200ps_crc32c_extend": (u32, [u32, vp, u64]),
    "b200ps_libsvm_parse": (i64, [C.c_char_p, u64, vp, vp, vp, vp, u64, u64]),
}
