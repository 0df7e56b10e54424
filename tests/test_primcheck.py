"""Arithmetic shortcuts of the CUDA stage code (plain arithmetic where the reference's saturating operators provably
cannot trigger, closed forms of operator chains) against the operator-by-operator forms: tests/hostsim/primcheck.cpp,
exhaustive where the domain is small, edge values + 2 x 10^7 random inputs elsewhere; and the no-overflow bound of the
FFT network (which lengths may run on plain additions)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_shortcuts_equal_operator_forms():
    d = os.path.join(HERE, "hostsim")
    subprocess.run(["make", "-s", "primcheck"], cwd=d, check=True)
    r = subprocess.run([os.path.join(d, "primcheck")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    for name in ("quant_bin", "m_log2_poly", "l_round", "l_shl_plan", "fft_bound", "normcorr_q15"):
        assert f"ok {name}" in r.stdout
