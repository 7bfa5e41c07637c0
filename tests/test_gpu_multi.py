"""GPU tests of the single-process multi-device handle (csrc/rt_multi.cu, include/rt_cuda.h rt_multi_*).

The reference's drivers construct ONE rt::Solver in ONE process (src/main.cc:100-113), so the group sharding for several
GPUs lives under the C ABI; the drop-in takes the device list from RT_CUDA_DEVICES.  Energy groups never couple in the
reference's solve, hence a sharded solve must reproduce the one-handle solve BIT FOR BIT; only the
temperature extension exchanges data (ncclAllReduce inside the library).  On a one-GPU box the sharding, the threads and
the gathers are exercised with device 0 listed several times; the cases that need two GPUs skip."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import PRM_DIR, ROOT
from parity_util import CELL_FLOOR, make_oracle, state_ok, synthetic_parameters

pytestmark = pytest.mark.gpu


def _device_count():
    import torch
    return torch.cuda.device_count()


def _all_results(s):
    s.compute_balance()
    s.compute_group_ends()
    return dict(psi=s.psi_mat_ref.copy(order="F"), ends=s.ends(0), prev=s.ends(1), mom=s.moments(), bal=s.get_balance(),
                left=s.get_ends("left"), right=s.get_ends("right"), B=s.group_terms()["B"])


def _assert_same(a, b):
    for k in ("psi", "ends", "prev", "bal", "left", "right", "B"):
        assert np.array_equal(a[k], b[k]), k
    for x, y in zip(a["mom"], b["mom"]):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("devices", [[0], [0, 0], [0, 0, 0]])
@pytest.mark.parametrize("bcl", [2, 1])
def test_shards_on_one_gpu_reproduce_the_single_handle_bitwise(rt, devices, bcl):
    p = synthetic_parameters(rt, M=8, G=7, N=3000, dt=1e-3, V=1.0, steps=2, bc_left=bcl, bc_right=1)  # 7 groups: uneven blocks
    with rt.Solver(p) as one:
        one.solve()
        want = _all_results(one)
    with rt.Solver(p, devices=devices) as s:
        sh = s.shards()
        assert len(sh) == len(devices) and sum(n for _, n, _ in sh) == 7 and [d for _, _, d in sh] == devices
        assert sh[0][0] == 0 and all(sh[k + 1][0] == sh[k][0] + sh[k][1] for k in range(len(sh) - 1))
        s.solve()
        assert s.stages_done == s.total_stages == 8
        _assert_same(_all_results(s), want)


def test_a_group_subrange_can_be_sharded(rt):
    """the multi handle owns [group_begin, group_begin + group_count) of a larger problem"""
    p = synthetic_parameters(rt, M=4, G=9, N=2500, dt=1e-3, steps=1)
    with rt.Solver(p) as whole:
        whole.solve()
        phi = whole.moments()[0]
    with rt.Solver(p, group_begin=2, group_count=5, devices=[0, 0]) as s:
        assert [(a, n) for a, n, _ in s.shards()] == [(2, 3), (5, 2)]
        s.solve()
        assert np.array_equal(s.moments()[0], phi[2:7])


def test_temperature_update_needs_distinct_devices(rt):
    p = synthetic_parameters(rt, M=4, G=6, N=500, dt=1e-3, steps=1)
    p.cv = 0.7
    with pytest.raises(rt.capi.RtError, match="DISTINCT device"):
        rt.Solver(p, devices=[0, 0])


def test_errors_of_a_shard_reach_the_caller(rt):
    p = synthetic_parameters(rt, M=4, G=3, N=500, dt=1e-3, steps=1)
    with pytest.raises(rt.capi.RtError, match="cannot be sharded over 4"):
        rt.Solver(p, devices=[0, 0, 0, 0])
    with pytest.raises(rt.capi.RtError, match="shard 1 on device 99"):
        rt.Solver(p, devices=[0, 99])


def test_cpp_dropin_with_several_devices_writes_identical_files(tmp_path):
    """RT_CUDA_DEVICES=<list> transfer_b200 <prm>: the reference's unmodified driver on the drop-in, groups sharded inside
    the library -- the 8 CSV files equal those of the one-device run byte for byte"""
    exe = os.path.join(ROOT, "radiative-transfer_b200", "host", "build", "transfer_b200")
    if not os.path.exists(exe):
        pytest.fail(f"{exe} missing: __graft_entry__.build() makes it where the reference checkout is present")
    shutil.copytree(PRM_DIR, tmp_path / "prm")
    two = "0,1" if _device_count() >= 2 else "0,0"
    outs = {}
    for tag, devs in (("one", None), ("many", two), ("three", "0,0,0")):
        run = tmp_path / tag
        run.mkdir()
        env = dict(os.environ)
        env.pop("RT_CUDA_DEVICES", None)
        if devs:
            env["RT_CUDA_DEVICES"] = devs
        out = subprocess.run([exe, str(tmp_path / "prm" / "llnl_slab_test.prm")], cwd=run, capture_output=True, text=True, timeout=600, env=env)
        assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
        if devs:
            assert "energy groups sharded over" in out.stdout
        outs[tag] = {f: open(run / f, "rb").read() for f in sorted(os.listdir(run)) if f.endswith(".csv")}
    assert len(outs["one"]) == 8
    assert outs["many"] == outs["one"] and outs["three"] == outs["one"]


def test_two_gpus_reproduce_one_bitwise(rt):
    if _device_count() < 2:
        pytest.skip("needs 2 GPUs")
    p = synthetic_parameters(rt, M=8, G=6, N=5000, dt=1e-3, V=1.0, steps=2)
    with rt.Solver(p) as one:
        one.solve()
        want = _all_results(one)
    with rt.Solver(p, devices=[0, 1]) as s:
        s.solve()
        _assert_same(_all_results(s), want)


def test_two_gpus_temperature_update_through_nccl_in_one_process(rt, oracle):
    """the path's one collective, inside ONE process: ncclCommInitAll + ncclAllReduce on the library's streams, one host
    thread per device; against the CPU oracle of the whole problem after every step (PARITY UNPINNED: the reference has
    no temperature update)"""
    if _device_count() < 2:
        pytest.skip("needs 2 GPUs")
    N, steps = 2000, 3
    p = synthetic_parameters(rt, M=8, G=12, N=N, dt=1e-3, V=1.0, steps=steps, bc_left=2, bc_right=1)
    x = (np.arange(N) + 0.5) / N
    p.T_cell = 0.8 + 0.4 * np.sin(2 * np.pi * x) ** 2
    p.rho_cell = 1.0 + 0.5 * np.cos(3 * np.pi * x)
    p.cv = 0.7
    o = make_oracle(oracle, rt, p, num_threads=4)
    with rt.Solver(p, devices=[0, 1]) as s:
        for _ in range(steps):
            o.run_stages(4)
            s.run_stages(4)
            T_ref, A, E, D = o.temperature(with_sums=True)
            T, a, e, d = s.temperature(with_sums=True)
            for got, ref in ((T, T_ref), (a, A), (e, E), (d, D)):
                assert np.max(np.abs(got - ref) / np.abs(ref)) <= 1e-9
            ok, worst, _ = state_ok(s.psi_mat_ref, o.psi(), np.maximum(o.group_terms()["B"], 1e-300), floor=CELL_FLOOR)
            assert ok, worst
        assert np.max(np.abs(s.temperature() - p.T_cell)) > 1e-4
    o.close()
