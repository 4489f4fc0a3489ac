"""ctypes binding of libngp_b200.so (the C ABI in include/ngp_b200.h).

There is no fallback: if the shared library is missing the import fails, and every compute
entry point fails with the library's own error when no CUDA device is present.
"""
import ctypes
import os

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG_DIR, "libngp_b200.so")

c_u32 = ctypes.c_uint32
c_u64 = ctypes.c_uint64
c_f32 = ctypes.c_float
c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t
c_int = ctypes.c_int
c_cp = ctypes.c_char_p


class NgpbError(RuntimeError):
    """Raised with the message the C ABI stored (same text as the reference's std::runtime_error)."""


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python instant-ngp-rendering_b200/build.py` "
            "(there is no CPU or PyTorch fallback)")
    return ctypes.CDLL(LIB_PATH)


lib = _load()

# name -> (restype, argtypes). Kept in one table so tests can check it against the header.
SIGNATURES = {
    "ngpb_last_error": (c_cp, []),
    "ngpb_version": (c_int, []),
    "ngpb_has_device": (c_int, []),
    "ngpb_batch_size_granularity": (c_u32, []),
    "ngpb_default_loss_scale": (c_f32, [c_int]),
    "ngpb_create_encoding": (c_int, [c_u32, c_cp, c_int, ctypes.POINTER(c_vp)]),
    "ngpb_create_network_with_input_encoding": (c_int, [c_u32, c_u32, c_cp, c_cp, ctypes.POINTER(c_vp)]),
    "ngpb_create_network": (c_int, [c_u32, c_u32, c_cp, ctypes.POINTER(c_vp)]),
    "ngpb_create_nerf_network": (c_int, [c_u32, c_u32, c_u32, c_u32, c_cp, c_cp, c_cp, c_cp, ctypes.POINTER(c_vp)]),
    "ngpb_module_free": (None, [c_vp]),
    "ngpb_module_n_input_dims": (c_u32, [c_vp]),
    "ngpb_module_n_output_dims": (c_u32, [c_vp]),
    "ngpb_module_n_output_dims_unpadded": (c_u32, [c_vp]),
    "ngpb_module_n_params": (c_sz, [c_vp]),
    "ngpb_module_param_precision": (c_int, [c_vp]),
    "ngpb_module_output_precision": (c_int, [c_vp]),
    "ngpb_module_hyperparams": (c_sz, [c_vp, c_cp, c_sz]),
    "ngpb_module_name": (c_sz, [c_vp, c_cp, c_sz]),
    "ngpb_module_layer_sizes": (c_sz, [c_vp, ctypes.POINTER(c_u32), c_sz]),
    "ngpb_module_initialize_params": (c_int, [c_vp, c_vp, c_u64, c_vp, c_f32]),
    "ngpb_module_inference": (c_int, [c_vp, c_vp, c_u32, c_vp, c_vp, c_vp]),
    "ngpb_static_scratch_bytes": (c_sz, []),
    "ngpb_module_inference_static_scratch": (c_int, [c_vp, c_vp, c_u32, c_vp, c_vp, c_vp]),
    "ngpb_module_forward": (c_int, [c_vp, c_vp, c_u32, c_vp, c_vp, c_vp, c_int, ctypes.POINTER(c_vp)]),
    "ngpb_module_backward": (c_int, [c_vp, c_vp, c_vp, c_u32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "ngpb_context_free": (None, [c_vp]),
    "ngpb_grid_level_table": (c_u32, [c_vp, ctypes.POINTER(c_u32), ctypes.POINTER(c_f32), ctypes.POINTER(c_u32)]),
    "ngpb_grid_indices": (c_int, [c_vp, c_vp, c_u32, c_vp, c_vp]),
    "ngpb_grid_set_max_level": (c_int, [c_vp, c_f32]),
    "ngpb_optimizer_create": (c_int, [c_cp, ctypes.POINTER(c_vp)]),
    "ngpb_optimizer_free": (None, [c_vp]),
    "ngpb_optimizer_allocate": (c_int, [c_vp, c_u32, ctypes.POINTER(c_u32), c_sz]),
    "ngpb_optimizer_step": (c_int, [c_vp, c_vp, c_f32, c_vp, c_vp, c_vp]),
    "ngpb_optimizer_learning_rate": (c_f32, [c_vp]),
    "ngpb_optimizer_set_learning_rate": (c_int, [c_vp, c_f32]),
    "ngpb_optimizer_step_count": (c_u32, [c_vp]),
    "ngpb_optimizer_custom_weights": (c_vp, [c_vp]),
    "ngpb_optimizer_update_hyperparams": (c_int, [c_vp, c_cp]),
    "ngpb_optimizer_hyperparams": (c_sz, [c_vp, c_cp, c_sz]),
    "ngpb_optimizer_state": (c_int, [c_vp, ctypes.POINTER(c_vp), ctypes.POINTER(c_vp), ctypes.POINTER(c_vp)]),
    "ngpb_nerf_network_density": (c_int, [c_vp, c_vp, c_u32, c_vp, c_u32, c_vp, c_u32, c_vp]),
    "ngpb_nerf_network_param_layout": (c_int, [c_vp, ctypes.POINTER(c_sz)]),
    "ngpb_nerf_mark_untrained_density_grid": (c_int, [c_vp, c_u32, c_vp, c_u32, c_vp, c_int]),
    "ngpb_nerf_generate_grid_samples": (c_int, [c_vp, c_u32, ctypes.POINTER(c_u64), c_u32, ctypes.POINTER(c_f32), c_vp, c_vp, c_vp, c_u32, c_f32]),
    "ngpb_nerf_splat_grid_samples": (c_int, [c_vp, c_u32, c_vp, c_vp, c_vp, c_int]),
    "ngpb_nerf_ema_grid_samples": (c_int, [c_vp, c_u32, c_f32, c_vp, c_vp]),
    "ngpb_nerf_generate_training_samples": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(c_u64)]),
    "ngpb_nerf_compute_loss_and_compact": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(c_u64)]),
    "ngpb_nerf_update_bitfield": (c_int, [c_vp, c_u32, c_vp, c_vp, c_vp]),
    "ngpb_nerf_bitfield_bytes": (c_sz, []),
    "ngpb_profile_enable": (None, [c_int]),
    "ngpb_profile_collect": (c_sz, [c_cp, c_sz, c_int]),
    "ngpb_testbed_set_stream": (c_int, [c_vp, c_vp]),
    "ngpb_testbed_create": (c_int, [ctypes.POINTER(c_vp)]),
    "ngpb_testbed_free": (None, [c_vp]),
    "ngpb_testbed_create_empty_nerf_dataset": (c_int, [c_vp, c_u32, c_int, c_int]),
    "ngpb_testbed_set_nerf_image": (c_int, [c_vp, c_u32, c_vp, c_int, c_int, c_int]),
    "ngpb_testbed_set_nerf_camera": (c_int, [c_vp, c_u32, ctypes.POINTER(c_f32), c_f32, c_f32, c_f32, c_f32, c_int, ctypes.POINTER(c_f32)]),
    "ngpb_testbed_reload_network_from_json": (c_int, [c_vp, c_cp]),
    "ngpb_testbed_reset_network": (c_int, [c_vp, c_int]),
    "ngpb_testbed_reload_network_from_file": (c_int, [c_vp, c_cp]),
    "ngpb_testbed_load_training_data": (c_int, [c_vp, c_cp, c_u32]),
    "ngpb_testbed_load_file": (c_int, [c_vp, c_cp]),
    "ngpb_testbed_save_snapshot": (c_int, [c_vp, c_cp, c_int, c_int]),
    "ngpb_testbed_load_snapshot": (c_int, [c_vp, c_cp]),
    "ngpb_testbed_image_metadata": (c_int, [c_vp, c_u32, c_vp]),
    "ngpb_testbed_view_state": (c_int, [c_vp, ctypes.POINTER(c_f32), c_int]),
    "ngpb_testbed_dataset_info": (c_sz, [c_vp, c_vp, c_sz]),
    "ngpb_testbed_set_dataset_info": (c_int, [c_vp, c_cp]),
    "ngpb_host_load_network_config": (c_sz, [c_cp, c_vp, c_sz]),
    "ngpb_host_describe_dataset": (c_sz, [c_cp, c_vp, c_sz]),
    "ngpb_host_msgpack_reencode": (c_sz, [c_vp, c_sz, c_vp, c_sz]),
    "ngpb_host_decode_png": (c_int, [c_cp, c_vp, c_sz, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "ngpb_host_natural_less": (c_int, [c_cp, c_cp]),
    "ngpb_host_decode_image": (c_int, [c_cp, c_vp, c_sz, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "ngpb_host_decode_hdr": (c_int, [c_cp, c_vp, c_sz, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "ngpb_host_decode_exr_frame": (c_int, [c_cp, c_vp, c_sz, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "ngpb_host_decode_exr": (c_int, [c_cp, c_vp, c_sz, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "ngpb_testbed_train": (c_int, [c_vp, c_u32]),
    "ngpb_testbed_get": (c_int, [c_vp, c_cp, ctypes.POINTER(ctypes.c_double)]),
    "ngpb_testbed_set": (c_int, [c_vp, c_cp, ctypes.c_double]),
    "ngpb_testbed_device_ptr": (c_int, [c_vp, c_cp, ctypes.POINTER(c_vp), ctypes.POINTER(c_sz)]),
    "ngpb_testbed_network_config": (c_sz, [c_vp, c_cp, c_sz]),
    "ngpb_testbed_update_density_grid": (c_int, [c_vp, c_f32, c_u32, c_u32]),
    "ngpb_probe_resident": (c_int, [c_vp, c_u32, c_u32, ctypes.c_uint64]),
    "ngpb_probe_gather": (c_int, [c_vp, c_vp, c_u32, c_u32, c_u32, c_u32, c_vp]),
    "ngpb_probe_scatter": (c_int, [c_vp, c_vp, c_u32, c_u32, c_u32]),
    "ngpb_probe_scatter_vec": (c_int, [c_vp, c_vp, c_u32, c_u32, c_u32, c_u32]),
    "ngpb_probe_l2_stream": (c_int, [c_vp, c_vp, c_sz, c_u32, c_u32, c_vp]),
    "ngpb_testbed_set_image": (c_int, [c_vp, c_vp, c_int, c_int, c_int]),
    "ngpb_testbed_render_image": (c_int, [c_vp, c_u32, c_u32, c_vp]),
    "ngpb_testbed_train_begin": (c_int, [c_vp, c_u32]),
    "ngpb_testbed_train_end": (c_int, [c_vp, c_u32]),
    "ngpb_testbed_set_render_lens": (c_int, [c_vp, c_int, ctypes.POINTER(c_f32)]),
    "ngpb_testbed_overlay_ground_truth": (c_int, [c_vp, c_u32, c_u32, c_int, c_u32, c_int, c_f32, c_f32, c_vp]),
    "ngpb_testbed_render": (c_int, [c_vp, c_u32, c_u32, c_u32, c_int, ctypes.POINTER(c_f32), c_f32, c_f32, c_f32, c_f32, c_vp]),
    "ngpb_testbed_train_forward": (c_int, [c_vp, c_u32]),
    "ngpb_testbed_set_params": (c_int, [c_vp, c_vp, c_int, c_sz]),
    "ngpb_testbed_set_density_grid_fp16": (c_int, [c_vp, c_vp, c_sz]),
    "ngpb_testbed_restore_training_state": (c_int, [c_vp, c_u32, c_f32]),
    "ngpb_testbed_train_backward": (c_int, [c_vp, c_u32]),
    "ngpb_testbed_set_render_steps_out": (c_int, [c_vp, c_vp]),
    "ngpb_testbed_render_nerf": (c_int, [c_vp, c_u32, c_u32, ctypes.POINTER(c_f32), c_f32, c_f32, c_f32, c_f32, c_u32, c_u32, c_u32, c_vp, c_vp]),
    "ngpb_selftest_umma": (c_int, [c_vp, c_int, c_u32, c_u32, c_vp, c_vp, c_vp]),
    "ngpb_nccl_get_unique_id": (c_int, [c_vp]),
    "ngpb_testbed_init_data_parallel": (c_int, [c_vp, c_vp, c_vp, c_int, c_int]),
    "ngpb_probe_umma_rate": (c_int, [c_vp, c_int, c_u32, c_u32, c_u32, c_vp]),
    "ngpb_random_uniform": (c_int, [c_vp, ctypes.POINTER(c_u64), c_sz, c_vp, c_f32, c_f32]),
}


def _bind():
    missing = []
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            missing.append(name)
            continue
        fn.restype = res
        fn.argtypes = args
    return missing


MISSING_SYMBOLS = _bind()


def check(status):
    if status != 0:
        raise NgpbError(lib.ngpb_last_error().decode("utf-8", "replace"))


def has_device():
    return bool(lib.ngpb_has_device())
