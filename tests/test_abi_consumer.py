"""The C ABI from plain C: tests/abi_consumer.c is compiled with gcc against include/mpc_b200.h + libmpcb200.so.
CPU part: the header is valid C99, the program links, and without a device mpc_create fails loudly (MPC_ECUDA).
GPU part: the C program's MPPI solve equals the oracle on the identical (LCG-generated) noise tensor."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "benchmarking-mpc-solvers_b200")


def _build(tmp_path):
    import mpc_b200
    mpc_b200.build_library()
    exe = str(tmp_path / "abi_consumer")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "abi_consumer.c"), "-L", LIBDIR, "-l:libmpcb200.so", "-Wl,-rpath," + LIBDIR,
           "-lm", "-o", exe]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return exe


@pytest.mark.skipif(shutil.which("gcc") is None, reason="gcc not available")
def test_header_is_c99_and_consumer_links_and_fails_loudly_without_a_device(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe, "probe"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("abi 5")
    assert "device present" in r.stdout or "no CPU path" in r.stdout or "create failed: -2" in r.stdout


def _lcg_normals(seed, n):
    s = np.uint64(seed)
    a, c = np.uint64(6364136223846793005), np.uint64(1442695040888963407)
    out = np.empty(n)
    with np.errstate(over="ignore"):
        for i in range(n):
            s = s * a + c
            u1 = (float(s >> np.uint64(11)) + 0.5) / 9007199254740992.0
            s = s * a + c
            u2 = (float(s >> np.uint64(11)) + 0.5) / 9007199254740992.0
            out[i] = np.sqrt(-2.0 * np.log(u1)) * np.cos(6.283185307179586 * u2)
    return out


@pytest.mark.gpu
def test_c_consumer_solve_equals_oracle(tmp_path):
    from oracle import oracle as orc
    exe = _build(tmp_path)
    K, H, seed = 500, 20, 12345
    r = subprocess.run([exe, "solve", str(K), str(H), str(seed)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = dict(l.split(" ", 1) for l in r.stdout.strip().split("\n"))
    eps = _lcg_normals(seed, K * H).reshape(K, H)
    e0 = [float(v) for v in lines["eps0"].split()]
    assert abs(eps[0, 0] - e0[0]) < 1e-12 and abs(eps[-1, -1] - e0[1]) < 1e-12  # same tensor on both sides
    m = orc.model("pendulum")
    x = np.array([3.0, -0.5])
    S, U_ref, u0_ref = orc.mppi_solve(m, x, np.zeros(H), eps, 0.5, 1.0)
    U = np.array([float(v) for v in lines["U"].split()])
    np.testing.assert_allclose(U, U_ref, rtol=1e-10, atol=1e-12)
    assert float(lines["u0"]) == pytest.approx(float(np.ravel(u0_ref)[0]), rel=1e-10)
    _, c_ref = orc.simulate(m, x, np.clip(U, -2.0, 2.0))
    assert float(lines["cost"]) == pytest.approx(c_ref, rel=1e-10)
