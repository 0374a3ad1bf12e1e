"""Reference-compatible command-line programs: stdout formats of SURVEY.md section 3.

The reference prints (riemann.cpp:92,94-96; cintegrate.cu:140-141; 4main.c:73,239,241)
    "%lf seconds"
    "The integral of f(x) from 0.0 to 3.14159265358979 with 1000000000 steps is <value>"
    "final distance is:%lf"
    "Step size of 10000" / "Total distance traveled = %lf"
Values are matched to the reference's own reproducibility (~1e-9 relative), its digits beyond
that are summation noise that changes with its rank count.
"""
import json
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "bin")


def run(name, *args):
    exe = os.path.join(BIN, name)
    if not os.path.exists(exe):
        import build
        build.build_cli()
    p = subprocess.run([exe] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                       timeout=300)
    assert p.returncode == 0, p.stderr
    return p.stdout.strip().splitlines()


def test_riemann_zero_arg_format_cpu():
    out = run("riemann", "--backend", "cpu")
    assert re.fullmatch(r"\d+\.\d{6} seconds", out[0])
    m = re.fullmatch(r"The integral of f\(x\) from 0\.0 to 3\.14159265358979 with 1000000000 steps is (\S+)",
                     out[1])
    assert m, out[1]
    # reference: 2.00000000000222 (1 worker) ... 2.0000000000002 (7 workers); exact left sum: 2 - 1.6e-18
    assert abs(float(m.group(1)) - 2.0) < 1e-12


def test_riemann_flags_and_json():
    out = run("riemann", "--backend", "cpu", "--n", "2^20", "--rule", "mid", "--json")
    j = json.loads(out[2])
    assert j["n"] == 1 << 20 and j["backend"] == "cpu" and j["rel_err_vs_oracle"] < 1e-15
    assert "with 1048576 steps" in out[1]


def test_cintegrate_zero_arg_cpu():
    out = run("cintegrate", "--backend", "cpu")
    assert re.fullmatch(r"\d+\.\d{6} seconds", out[0])
    assert out[1] == "final distance is:122000.004000"


def test_cintegrate_compat_reproduces_reference_coverage():
    # 1800/64 truncates to 28 s per thread in the reference: 1792 of 1800 s -> 121999.800663
    out = run("cintegrate", "--backend", "cpu", "--compat")
    assert out[1] == "final distance is:121999.800663"


def test_cintegrate_sin_kernel():
    out = run("cintegrate", "--backend", "cpu", "--sin", "--n", "1000000")
    assert out[1] == "final distance is:2.000000"


def test_train4_zero_arg_cpu():
    out = run("train4", "--backend", "cpu", "--json")
    assert out[0] == "Step size of 10000"
    assert re.fullmatch(r"\d+\.\d{6} seconds", out[1])
    # reference prints 122000.004030 (P=1) .. 122000.003997 (P=8); exactly rounded: 122000.004000
    assert out[2] == "Total distance traveled = 122000.004000"
    j = json.loads(out[3])
    assert abs(j["q_last"] / 1e8 - 1.0986e8) / 1.0986e8 < 1e-4


@pytest.mark.gpu
def test_clis_on_gpu():
    out = run("riemann", "--world", "1", "--json")
    assert abs(json.loads(out[2])["value"] - 2.0) < 1e-12 and json.loads(out[2])["backend"] == "cuda"
    assert run("cintegrate", "--world", "1")[1] == "final distance is:122000.004000"
    assert run("train4", "--world", "1")[2] == "Total distance traveled = 122000.004000"
