"""ctypes binding of lib/libgeotx_b200.so -- the C ABI declared in include/geotx_b200.h.

There is no fallback of any kind: if the library is missing or a call fails, a RuntimeError is raised
(the reference's own extension reports errors the same way, TORCH_CHECK -> RuntimeError,
/root/reference/geotransformer/extensions/common/torch_helper.h:6-35)."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('GEOTX_LIB') or os.path.join(_HERE, 'lib', 'libgeotx_b200.so')  # GEOTX_LIB: timing experiments

_p = ctypes.c_void_p
_i = ctypes.c_int
_f = ctypes.c_float
_i64 = ctypes.c_int64

# name -> (restype, argtypes); must list every symbol include/geotx_b200.h declares (tests check that)
SIGNATURES = {
    'geotx_last_error': (ctypes.c_char_p, []),
    'geotx_abi_version': (_i, []),
    'geotx_copy_async': (_i, [_p, _p, _i64, _p]),
    'geotx_gse_geometry_bytes': (_i64, [_i, _i, _i]),
    'geotx_gse_knn': (_i, [_p, _i, _i, _i, _p, _p, _p]),
    'geotx_gse_indices': (_i, [_p, _i, _i, _i, _f, _f, _p, _p, _p]),
    'geotx_gse_packed_weight_bytes': (_i64, [_i]),
    'geotx_gse_pack_weights': (_i, [_p, _p, _p, _p, _i, _p, _p]),
    'geotx_gse_embed': (_i, [_p, _i, _i, _i, _i, _i, _p, _p, _i, _f, _f, _i, _i, _p, _p]),
    'geotx_gse_embed_scores': (_i, [_p, _i, _i, _i, _i, _i, _p, _p, _i, _f, _f, _i, _i, _p, _p, _i64, _i64, _i64, _i, _p, _p]),
    'geotx_watchdog_status': (_i, [_p]),
    'geotx_gse_embed_fp32': (_i, [_p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _i, _f, _f, _i, _i, _p, _p]),
    'geotx_rpe_scores': (_i, [_p, _i64, _i64, _i64, _p, _i, _i, _i, _i, _i, _i, _i, _i, _p, _p]),
    'geotx_attention': (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _f, _p, _i, _p]),
    'geotx_attention_f16': (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _f, _p, _i, _p]),
    'geotx_add_layernorm': (_i, [_p, _p, _p, _p, _i64, _i, _f, _p, _p]),
    'geotx_spm_workspace_bytes': (_i64, [_i, _i, _i, _i]),
    'geotx_superpoint_matching': (_i, [_p, _p, _i, _i, _i, _p, _p, _i, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p]),
    'geotx_p2n_workspace_bytes': (_i64, [_i, _i, _i]),
    'geotx_point_to_node_partition': (_i, [_p, _i, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p]),
    'geotx_p2n_static_scratch_bytes': (_i64, []),
    'geotx_point_to_node_partition_static': (_i, [_p, _i, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p]),
    'geotx_log_optimal_transport': (_i, [_p, _p, _p, _p, _i, _f, _i, _i, _i, _p, _p]),
    'geotx_weighted_procrustes': (_i, [_p, _p, _p, _i, _i, _f, _f, _p, _p]),
    'geotx_radius_neighbors': (_i, [_p, _p, _p, _p, _i, _f, _i, _p, _p, _p, _p]),
    'geotx_radius_neighbors_workspace_bytes': (_i64, [_i64, _i64]),
    'geotx_radius_neighbors_grid': (_i, [_p, _p, _p, _p, _i, _f, _i, _p, _p, _p, _p, _i64, _i64, _p]),
    'geotx_grid_subsample_workspace_bytes': (_i64, [_i64]),
    'geotx_grid_subsample': (_i, [_p, _p, _i, _f, _p, _p, _p, _i64, _p]),
    'geotx_lgr_workspace_bytes': (_i64, [_i, _i]),
    'geotx_local_global_registration': (_i, [_p, _p, _p, _p, _p, _i64, _i64, _p, _i, _i, _i, _f, _i, _f, _i, _i,
                                             _p, _p, _p, _p, _p, _p, _p]),
    'geotx_gse_table_bytes': (_i64, [_i, _i, _i]),
    'geotx_gse_build_tables': (_i, [_p, _p, _p, _p, _p, _i, _f, _f, _i, _i, _p, _p]),
    'geotx_gse_embed_table': (_i, [_p, _i, _i, _i, _i, _i, _p, _i, _i, _i, _f, _f, _f, _f, _p, _p, _p, _p, _p, _p]),
    'geotx_gse_cheb_ka_tiles': (_i, [_f]),
    'geotx_gse_cheb_d_range': (_f, [_i]),
    'geotx_gse_cheb_bytes': (_i64, [_i, _i]),
    'geotx_gse_cheb_plain_rows': (_i64, [_i]),
    'geotx_gse_build_cheb': (_i, [_p, _p, _p, _p, _p, _i, _f, _i, _p, _p, _p, _p]),
    'geotx_gse_embed_cheb': (_i, [_p, _i, _i, _i, _i, _i, _p, _i, _f, _i, _i, _f, _f, _i, _i, _p, _p, _p]),
    'geotx_rpe_scores_cheb': (_i, [_p, _i, _i, _i, _i, _i, _p, _i, _f, _i, _i, _f, _f, _i, _p, _i64, _i64, _i64, _i,
                                    _p, _i64, _i64, _i64, _p, _p, _p]),
    'geotx_f32_to_f16': (_i, [_p, _p, _i64, _p]),
    'geotx_pack_linear_weight': (_i, [_p, _i, _i, _p, _p]),
    'geotx_linear_f16': (_i, [_p, _i64, _i64, _i, _p, _p, _i, _i, _p, _i64, _p]),
    'geotx_linear_f16_mixed': (_i, [_p, _i64, _i64, _i, _p, _p, _i, _i, _p, _i64, _i, _p, _i64, _p]),
    'geotx_post_attention': (_i, [_p, _i64, _p, _i64, _i64, _i, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _p, _f, _p, _i64,
                                  _p]),
    'geotx_mma_sync_probe': (_i, [_i, _i, _p, _p]),
    'geotx_patch_scores': (_i, [_p, _i64, _p, _i64, _i, _p, _p, _i, _p, _p, _i, _p, _p]),
}

GEOTX_F32 = 0
GEOTX_BF16 = 1
GEOTX_F16 = 2

# kernels each entry point enqueues (bench.py's gpu_launches is counted from this)
KERNELS_PER_CALL = {'geotx_copy_async': 0,  # copy engines, no kernel
                    'geotx_gse_knn': 2,  # pack_points + knn
                    'geotx_superpoint_matching': 10,  # prepare, scores, sums, 6 select passes, collect
                    'geotx_point_to_node_partition': 4,  # assign, offsets, scatter, select
                    'geotx_point_to_node_partition_static': 4,
                    'geotx_grid_subsample': 14,
                    'geotx_local_global_registration': 4}  # 7 per batch element (bounds, clear, accumulate, count, scan, emit, advance)

_lib = None
launch_count = 0  # hand-written kernels launched through this binding so far


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f'{LIB_PATH} not found: the CUDA library has not been built (run `python -c "import '
                f'__graft_entry__ as g; g.build()"`).  ref-geotransformer_b200 has no CPU / PyTorch fallback.')
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def watchdog():
    """Reads and clears the tcgen05 kernel's protocol watchdog (synchronises).  Returns None when healthy."""
    rec = (ctypes.c_uint32 * 4)()
    check(load().geotx_watchdog_status(rec), 'geotx_watchdog_status')
    if rec[0] == 0:
        return None
    roles = {1: 'producer<-empty', 2: 'epilogue<-acc_full', 3: 'mma<-acc_empty', 4: 'mma<-full', 5: 'loader<-empty'}
    return dict(role=roles.get((rec[0] >> 24) & 0x7f, '?'), ring_pos=rec[0] & 0xFFFFFF, barrier=hex(rec[1]),
                parity=rec[2], block=rec[3])


def check(rc, what):
    if rc != 0:
        msg = load().geotx_last_error().decode('utf-8', 'replace')
        raise RuntimeError(f'{what} failed (code {rc}): {msg}')


def call(name, *args):
    global launch_count
    lib = load()
    rc = getattr(lib, name)(*args)
    check(rc, name)
    launch_count += KERNELS_PER_CALL.get(name, 1)
