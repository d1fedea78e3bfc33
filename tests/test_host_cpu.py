"""CPU: host-side logic of the binding that needs no device."""
import copy
import os

import pytest


def _cfg(prop, **kw):
    from gapflow_b200 import _capi
    from tests.cases import inputs
    d, _ = inputs('journal1d_readme')
    p = copy.deepcopy(d['properties'])
    p.update(prop)
    try:
        return _capi.make_cfg(d['grid'], d['geometry'], p, d['numerics'], **kw), p
    except _capi.GapflowLibraryError:
        pytest.skip("library not built")


def test_viscosity_blocks_select_the_field_path():
    """A 'piezo' / 'thinning' block (stress.py:305-326) makes the viscosity a field: cfg.visc_field is set
    (never silently Newtonian) and the law ids / parameter order are those of include/gapflow_b200.h."""
    from gapflow_b200 import _capi
    c, _ = _cfg({})
    assert c.visc_field == 0
    c, p = _cfg({'piezo': {'name': 'Barus', 'aB': 2e-8}})
    assert c.visc_field == 1
    pz, pp, th, tp = _capi.viscosity_args(p)
    assert (pz, list(pp), th, tp) == (2, [2e-8], 0, None)
    c, p = _cfg({'thinning': {'name': 'Carreau', 'mu_inf': 1e-3, 'lam': 0.02, 'a': 2., 'N': 0.8},
                 'piezo': {'name': 'Roelands', 'mu_inf': 1e-3, 'p_ref': 1.96e8, 'z': 0.68}})
    assert c.visc_field == 1
    pz, pp, th, tp = _capi.viscosity_args(p)
    assert (pz, list(pp), th, list(tp)) == (3, [1e-3, 1.96e8, 0.68], 3, [1e-3, 0.02, 2., 0.8])
    # a name the reference does not know: identity law, still the field path (viscosity.py:64, 94)
    c, p = _cfg({'piezo': {'name': 'none'}, 'thinning': {'name': 'Eyring', 'tauE': 5e5}})
    assert _capi.viscosity_args(p)[0] == 1 and _capi.viscosity_args(p)[2] == 2 and c.visc_field == 1
    for name, want in (('Dukler', 4), ('McAdams', 5)):
        _, p = _cfg({'piezo': {'name': name, 'eta_v': 3.9e-5, 'rho_l': 850., 'rho_v': 0.019}})
        pz, pp, _, _ = _capi.viscosity_args(p)
        assert pz == want and list(pp) == [3.9e-5, 850., 0.019]


def test_viscosity_blocks_are_refused_on_slabs():
    """The pressure gradient at a halo row needs a second halo row: refused, not silently wrong."""
    with pytest.raises(NotImplementedError):
        _cfg({'thinning': {'name': 'Eyring', 'tauE': 5e5}}, halo=(True, False))


def test_bench_prints_its_line_once_whatever_slab_parity_does(tmp_path):
    """bench.py at N > 1: slab_parity runs under a deadline after the measurement.  A hang or an exception in it must
    still leave exactly ONE JSON line on stdout and exit code 0 (os._exit from the watchdog); a normal return too."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "deadline.py"
    script.write_text(
        "import sys, time\n"
        f"sys.path.insert(0, {root!r})\n"
        "import bench\n"
        "mode = sys.argv[1]\n"
        "line = {'metric': 'x', 'slab_parity': None}\n"
        "def finish(par):\n"
        "    line['slab_parity'] = par\n"
        "    bench.emit(line)\n"
        "def fn():\n"
        "    if mode == 'hang': time.sleep(60)\n"
        "    if mode == 'raise': raise RuntimeError('boom')\n"
        "    return {'bit_exact': True}\n"
        "finish(bench.run_with_deadline(fn, 1.0, 0, finish))\n"
        "finish({'bit_exact': 'second call must be ignored'})\n")
    for mode, want in (('ok', True), ('hang', None), ('raise', None)):
        r = subprocess.run([sys.executable, str(script), mode], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, (mode, r.stderr[-500:])
        lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
        assert len(lines) == 1, (mode, lines)
        par = json.loads(lines[0])['slab_parity']
        assert par['bit_exact'] is want, (mode, par)
        if mode == 'hang':
            assert 'did not finish' in par['error']
        if mode == 'raise':
            assert 'boom' in par['error']


def _dryrun(tmp_path, hang, extra):
    """bench.py at N = 2 on the CPU: gloo, stand-in solvers (tests/bench_dryrun.py).  -> [(exit code, stdout lines, stderr)]"""
    import socket
    import subprocess
    import sys
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        port = so.getsockname()[1]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        if hang:
            env["DRYRUN_HANG"] = hang
        procs.append(subprocess.Popen([sys.executable, os.path.join(root, "tests", "bench_dryrun.py"), "--gpus", "2", "--steps", "5",
                                       "--warmup", "3", "--cells", "64"] + extra, env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    out = []
    for p in procs:
        o, e = p.communicate(timeout=300)
        out.append((p.returncode, [ln for ln in o.splitlines() if ln.strip()], e))
    return out


def test_bench_two_ranks_dry_run(tmp_path):
    """The whole N > 1 control flow of bench.py (collectives in the same order on every rank, the JSON line, slab_parity
    after the measurement) with two gloo ranks; slab_parity's bit-equality here also says that bench.initial_fields gives
    the slabs the rows of the single domain's field."""
    import json
    (rc0, l0, e0), (rc1, l1, e1) = _dryrun(tmp_path, None, [])
    assert rc0 == 0 and rc1 == 0, (e0[-800:], e1[-800:])
    assert len(l0) == 1 and l1 == []
    d = json.loads(l0[0])
    assert d["n_gpus"] == 2 and d["steps"] == 5 and d["value"] > 0 and d["scaling"] == "weak"
    assert d["slab_parity"]["bit_exact"] is True and d["slab_parity"]["arith"] == ["fast", "exact"]
    for key in ("metric", "unit", "warmup", "ms_per_step", "higher_is_better", "vs_baseline", "dtype", "data", "config", "clocks",
                "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert key in d, key
    assert d["e2e"]["h2d_bytes_per_step"] == 2 * 3 * 66 * 66 * 8


def test_bench_two_ranks_with_a_peer_that_hangs(tmp_path):
    """A rank that hangs in slab_parity: the measured line is printed with the reason, every rank exits 0.  A rank that hangs
    in the measurement: rank 0 prints a line with value null and an error, exit code 1 -- bounded, never a silent hang."""
    import json
    (rc0, l0, e0), (rc1, l1, e1) = _dryrun(tmp_path, "parity", ["--parity-deadline", "4"])
    assert rc0 == 0 and rc1 == 0, (e0[-800:], e1[-800:])
    assert len(l0) == 1 and l1 == []
    d = json.loads(l0[0])
    assert d["value"] > 0 and d["slab_parity"]["bit_exact"] is None and "did not finish" in d["slab_parity"]["error"]
    (rc0, l0, e0), (rc1, l1, e1) = _dryrun(tmp_path, "main", ["--deadline", "4"])
    assert rc0 == 1 and rc1 == 1, (rc0, rc1, e0[-800:], e1[-800:])
    assert len(l0) == 1 and l1 == []
    d = json.loads(l0[0])
    assert d["value"] is None and d["n_gpus"] == 2 and "did not finish" in d["error"]
