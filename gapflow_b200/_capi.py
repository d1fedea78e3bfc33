"""ctypes binding of libgapflow_b200.so (include/gapflow_b200.h).

The library is the product: there is no Python/NumPy fallback.  Importing this module
builds nothing; ``load()`` raises ``GapflowLibraryError`` if the shared library is missing
(run ``python -m gapflow_b200.build``) or does not export every declared symbol.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgapflow_b200.so")


class GapflowLibraryError(RuntimeError):
    pass


class GapflowError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gapflow_b200 error {code}: {msg}")
        self.code = code


GPF_OK, GPF_E_INVALID, GPF_E_CUDA, GPF_E_NOMEM, GPF_E_STATE = 0, -1, -2, -3, -4
EOS_IDS = {'DH': 0, 'PL': 1, 'vdW': 2, 'MT': 3, 'cubic': 4, 'BWR': 5, 'Bayada': 6}
EOS_KEYS = {'DH': ['rho0', 'P0', 'C1', 'C2'], 'PL': ['rho0', 'P0', 'alpha'], 'vdW': ['M', 'T', 'a', 'b'],
            'MT': ['rho0', 'P0', 'K', 'n'], 'cubic': ['a', 'b', 'c', 'd'], 'BWR': ['T', 'gamma'],
            'Bayada': ['rho_l', 'rho_v', 'c_l', 'c_v']}
BC_P, BC_D, BC_N, BC_HALO = 0, 1, 2, 3
F_Q, F_TOPO, F_EXTRA = 0, 1, 2
PART_ALL, PART_EDGE, PART_INTERIOR = 0, 1, 2
IPC_HANDLE_BYTES = 64
PEER_BLOB_BYTES = 96
SLIP = {'top': 0, 'both': 1}
PIEZO_IDS = {'Barus': 2, 'Roelands': 3, 'Dukler': 4, 'McAdams': 5}       # GPF_PIEZO_*; any other name: 1 (identity)
PIEZO_KEYS = {'Barus': ['aB'], 'Roelands': ['mu_inf', 'p_ref', 'z'], 'Dukler': ['eta_v', 'rho_l', 'rho_v'],
              'McAdams': ['eta_v', 'rho_l', 'rho_v']}
THIN_IDS = {'Eyring': 2, 'Carreau': 3}                                    # GPF_THIN_*
THIN_KEYS = {'Eyring': ['tauE'], 'Carreau': ['mu_inf', 'lam', 'a', 'N']}
FIN_STEP, FIN_PRERUN, FIN_PRERUN_KEEP_EKIN, FIN_REFRESH = 0, 1, 2, 3


class FvCfg(C.Structure):
    _fields_ = [("Nx", C.c_int32), ("Ny", C.c_int32), ("dx", C.c_double), ("dy", C.c_double),
                ("bc_type", (C.c_int32 * 3) * 4), ("bc_src", (C.c_int32 * 3) * 4), ("bc_val", C.c_double * 4),
                ("U", C.c_double), ("V", C.c_double), ("eta", C.c_double), ("zeta", C.c_double),
                ("eos", C.c_int32), ("eos_par", C.c_double * 8), ("rho0", C.c_double),
                ("adaptive", C.c_int32), ("CFL", C.c_double), ("dt", C.c_double), ("tol", C.c_double),
                ("max_it", C.c_int32), ("mc_order", C.c_int32), ("has_extra", C.c_int32),
                ("halo_edge", C.c_int32 * 2), ("arith", C.c_int32), ("visc_field", C.c_int32),
                ("reserved", C.c_int32 * 3)]


class FvScalars(C.Structure):
    _fields_ = [("step", C.c_int64), ("simtime", C.c_double), ("dt", C.c_double), ("residual", C.c_double),
                ("residual_buffer", C.c_double * 5), ("residual_count", C.c_int32),
                ("ekin", C.c_double), ("ekin_old", C.c_double), ("v_max", C.c_double), ("v_sound", C.c_double),
                ("dt_crit", C.c_double), ("cfl", C.c_double), ("has_nan", C.c_int32), ("has_neg", C.c_int32),
                ("stop", C.c_int32), ("converged", C.c_int32)]


_vp, _dp, _sz, _i = C.c_void_p, C.c_void_p, C.c_size_t, C.c_int
_cfgp = C.POINTER(FvCfg)

# name -> (restype, argtypes); must list every symbol of include/gapflow_b200.h
SIGNATURES = {
    "gpf_last_error": (C.c_char_p, []),
    "gpf_version": (_i, []),
    "gpf_launch_count": (C.c_int64, []),
    "gpf_fv_cfg_set_bc": (_i, [_cfgp, C.POINTER(C.c_uint8), C.POINTER(C.c_double)]),
    "gpf_fv_workspace_bytes": (_i, [_cfgp, C.POINTER(C.c_size_t)]),
    "gpf_fv_create": (_i, [_cfgp, _vp, _sz, C.POINTER(_vp)]),
    "gpf_fv_destroy": (_i, [_vp]),
    "gpf_fv_upload": (_i, [_vp, _i, _dp, _i, _i, _vp]),
    "gpf_fv_download": (_i, [_vp, _i, _dp, _i, _i, _vp]),
    "gpf_fv_pre_run": (_i, [_vp, _i, _vp]),
    "gpf_fv_step": (_i, [_vp, _i, _i, _vp]),
    "gpf_fv_fill_ghosts": (_i, [_vp, _vp]),
    "gpf_fv_get_scalars": (_i, [_vp, C.POINTER(FvScalars), _vp]),
    "gpf_fv_set_scalars": (_i, [_vp, C.POINTER(FvScalars), _vp]),
    "gpf_fv_mass": (_i, [_vp, C.POINTER(C.c_double), _vp]),
    "gpf_fv_closures": (_i, [_vp, _dp, _dp, _dp, _dp, _i, _vp]),
    "gpf_fv_closures_of": (_i, [_vp, _i, _dp, _dp, _dp, _dp, _i, _vp]),
    "gpf_fv_set_viscosity": (_i, [_vp, _i, C.POINTER(C.c_double), _i, C.POINTER(C.c_double), _dp, _dp]),
    "gpf_fv_viscosity": (_i, [_vp, _i, _dp, _i, _vp]),
    "gpf_fv_update_host": (_i, [_vp, _dp, _i, C.POINTER(FvScalars), _vp]),
    "gpf_fv_set_run_mode": (_i, [_vp, _i]),
    "gpf_fv_pass": (_i, [_vp, _i, _vp]),
    "gpf_fv_pass_part": (_i, [_vp, _i, _i, _vp]),
    "gpf_comm_create": (_i, [_i, _i, _i, _vp, C.POINTER(C.c_void_p)]),
    "gpf_comm_connect": (_i, [_vp, _vp]),
    "gpf_comm_destroy": (_i, [_vp]),
    "gpf_comm_error": (_i, [_vp, C.POINTER(C.c_int)]),
    "gpf_comm_set_timeout": (_i, [_vp, C.c_double]),
    "gpf_fv_peer_export": (_i, [_vp, _vp]),
    "gpf_comm_attach": (_i, [_vp, _vp, _vp, _i, _i]),
    "gpf_comm_detach": (_i, [_vp]),
    "gpf_comm_connect_local": (_i, [_vp, C.POINTER(C.c_void_p)]),
    "gpf_comm_attach_local": (_i, [_vp, _vp, _vp, _i, _vp, _i]),
    "gpf_fv_slab_pass": (_i, [_vp, _vp, _i, _vp]),
    "gpf_fv_halo_send": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "gpf_fv_halo_recv": (_i, [_vp, _vp, _i, _vp]),
    "gpf_fv_record_send": (_i, [_vp, _vp, _i, _vp]),
    "gpf_fv_finish_recv": (_i, [_vp, _vp, _i, _vp, _vp]),
    "gpf_fv_slab_step": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "gpf_fv_halo_pack": (_i, [_vp, _i, _dp, _vp]),
    "gpf_fv_halo_unpack": (_i, [_vp, _i, _dp, _vp]),
    "gpf_fv_reduce_local": (_i, [_vp, _i, _dp, _vp]),
    "gpf_fv_finish": (_i, [_vp, _i, _dp, _i, _vp]),
    "gpf_fv_set_timing": (_i, [_vp, _i]),
    "gpf_fv_get_timing": (_i, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "gpf_eos_pressure": (_i, [_cfgp, _dp, _dp, _sz, _vp]),
    "gpf_eos_sound_velocity": (_i, [_cfgp, _dp, _dp, _sz, _vp]),
    "gpf_stress_bottom": (_i, [_cfgp, _dp, _dp, _dp, _dp, _sz, _vp]),
    "gpf_stress_top": (_i, [_cfgp, _dp, _dp, _dp, _dp, _sz, _vp]),
    "gpf_stress_avg": (_i, [_cfgp, _dp, _dp, _dp, _dp, _sz, _vp]),
    "gpf_stress_ex": (_i, [_cfgp, _i, _i, _dp, _dp, _dp, _dp, _dp, _dp, _sz, _vp]),
    "gpf_selftest_stress_ex_host": (_i, [_cfgp, _i, _i, _dp, _dp, _dp, _dp, _dp, _dp, _sz]),
    "gpf_gp_workspace_bytes": (_i, [_i, _i, _i, C.POINTER(C.c_size_t)]),
    "gpf_gp_create": (_i, [_i, _i, _i, _vp, _sz, C.POINTER(_vp)]),
    "gpf_gp_destroy": (_i, [_vp]),
    "gpf_gp_set_arith": (_i, [_vp, _i]),
    "gpf_gp_factor": (_i, [_vp, _dp, _i, C.c_double, _dp, C.c_double, _dp, _vp]),
    "gpf_gp_get_factor": (_i, [_vp, _dp, _dp, _vp]),
    "gpf_gp_train_workspace_bytes": (_i, [_i, C.POINTER(C.c_size_t)]),
    "gpf_gp_nlml": (_i, [_vp, _vp, C.c_size_t, _dp, _dp, _vp]),
    "gpf_gp_mean": (_i, [_vp, _dp, C.c_int64, _dp, C.POINTER(C.c_double), _vp]),
    "gpf_gp_mean_var": (_i, [_vp, _dp, C.c_int64, _dp, C.c_double, _dp, _dp, C.POINTER(C.c_double),
                             C.POINTER(C.c_int64), _vp]),
    "gpf_gp_features": (_i, [_vp, _dp, C.POINTER(C.c_int32), _i, _dp, _vp]),
    "gpf_gp_features_of": (_i, [_vp, _i, _dp, C.POINTER(C.c_int32), _i, _dp, _vp]),
    "gpf_fv_set_surrogate": (_i, [_vp, _dp, C.c_double, _dp, C.c_double, _dp, C.c_double]),
    "gpf_fv_finish_surrogate": (_i, [_vp, _i, _dp, _i, C.c_double, _vp]),
    "gpf_selftest_mean1d_host": (_i, [_dp, _i, C.c_double, C.c_double, _dp, _dp, C.c_int64, _dp, _dp]),
    "gpf_selftest_var1d_host": (_i, [_dp, _i, C.c_double, C.c_double, C.c_double, _dp, C.c_int64, C.c_double, _dp]),
    "gpf_selftest_division": (_i, [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), _vp]),
}

_lib = None


def load():
    """Load the shared library and bind every declared symbol (fails loudly)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GapflowLibraryError(
            f"{LIB_PATH} not found: the CUDA library is the only implementation of this path "
            "(no CPU fallback). Build it with `python -m gapflow_b200.build`.")
    try:
        lib = C.CDLL(LIB_PATH)
    except OSError as e:
        raise GapflowLibraryError(f"cannot load {LIB_PATH}: {e}") from e
    missing = []
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            missing.append(name)
            continue
        fn.restype = res
        fn.argtypes = args
    if missing:
        raise GapflowLibraryError(f"{LIB_PATH} lacks symbols: {missing}")
    _lib = lib
    return lib


def check(rc):
    if rc != GPF_OK:
        raise GapflowError(rc, load().gpf_last_error().decode())


ARITH = {'exact': 0, 'fast': 1}


def make_cfg(grid, geo, prop, numerics, has_extra=False, halo=(False, False), halo_edge=(False, False),
             arith='exact'):
    """Flatten the reference's sanitised dicts (GaPFlow/io.py:139-395) into gpf_fv_cfg.

    ``halo=(lo, hi)`` marks x sides whose ghost row comes from a neighbour rank (multi-GPU);
    ``halo_edge`` marks those of them that are the global domain's periodic edge."""
    lib = load()
    c = FvCfg()
    c.Nx, c.Ny, c.dx, c.dy = grid['Nx'], grid['Ny'], grid['dx'], grid['dy']
    # The reference reads the Dirichlet value of the OPPOSITE-named side whenever a side is not periodic, for Neumann
    # components too (problem.py:682-707 -> :738-753), and its sanitiser stores that value only for sides that have a
    # 'D' component (io.py:183-191, 212-220): on a non-periodic axis both sides need one, or the first ghost fill ends
    # in this KeyError.  Same inputs refused, same exception, here already at construction.
    for side, looked_up in (('xE', 'xW'), ('xW', 'xE'), ('yS', 'yN'), ('yN', 'yS')):
        if not all(grid[f'bc_{side}_P']) and f'bc_{looked_up}_D_val' not in grid:
            raise KeyError(f'bc_{looked_up}_D_val')
    masks = (C.c_uint8 * 36)()
    vals = (C.c_double * 4)()
    for s, side in enumerate(['xE', 'xW', 'yS', 'yN']):
        for k, kind in enumerate(['P', 'D', 'N']):
            for comp in range(3):
                masks[(s * 3 + k) * 3 + comp] = 1 if grid[f'bc_{side}_{kind}'][comp] else 0
        v = grid.get(f'bc_{side}_D_val', 0.0)
        vals[s] = 0.0 if v is None else float(v)
    check(lib.gpf_fv_cfg_set_bc(C.byref(c), masks, vals))
    for s, h in enumerate(halo):
        if h:
            for comp in range(3):
                c.bc_type[s][comp] = BC_HALO
            c.halo_edge[s] = 1 if halo_edge[s] else 0
    c.U, c.V = geo['U'], geo['V']
    c.eta, c.zeta = prop['shear'], prop['bulk']
    # stress.py:306-326, 430-450: a 'piezo' / 'thinning' block makes the viscosity a field (even with a law
    # name the reference does not know: identity law, same field path); gpf_fv_set_viscosity() must follow
    c.visc_field = 1 if ('piezo' in prop or 'thinning' in prop) else 0
    if c.visc_field and any(halo):
        raise NotImplementedError("properties.piezo / properties.thinning on a slab-decomposed grid: the pressure "
                                  "gradient at a halo row needs a second halo row (not built); run on one GPU")
    if prop['EOS'] not in EOS_IDS:
        raise GapflowError(GPF_E_INVALID, f"EOS {prop['EOS']!r} has no fixed-form branch")
    c.eos = EOS_IDS[prop['EOS']]
    defaults = {'gamma': 3.0}
    for k, key in enumerate(EOS_KEYS[prop['EOS']]):
        c.eos_par[k] = float(prop[key]) if key in prop else defaults[key]
    c.rho0 = prop['rho0']
    c.adaptive = 1 if numerics['adaptive'] else 0
    c.CFL, c.dt, c.tol = numerics['CFL'], numerics['dt'], numerics['tol']
    c.max_it, c.mc_order = numerics['max_it'], numerics['MC_order']
    c.has_extra = 1 if has_extra else 0
    c.arith = ARITH[arith]
    return c


def viscosity_args(prop):
    """(piezo id, piezo parameters, thinning id, thinning parameters) of gpf_fv_set_viscosity from the
    sanitised properties dict (GaPFlow/io.py:318-362); parameter order as include/gapflow_b200.h states."""
    def one(sec, ids, keys):
        if sec not in prop:
            return 0, None
        name = prop[sec].get('name', 'none')
        if name not in ids:
            return 1, None                  # unknown name: the reference applies the identity law
        vals = [float(prop[sec][k]) for k in keys[name]]
        return ids[name], (C.c_double * len(vals))(*vals)
    return one('piezo', PIEZO_IDS, PIEZO_KEYS) + one('thinning', THIN_IDS, THIN_KEYS)
