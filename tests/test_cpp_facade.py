"""The C++ facade include/cstone_b200/*.hpp (cstone::Domain<KeyType,T,GpuTag>, DeviceVector, Box): builds with plain g++
against the C ABI (CPU check) and passes the reference-shaped domain test on the GPU (tests/cpp/domain_facade_test.cpp,
after [ref: test/integration_mpi/domain_nranks.cpp])."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "domain_facade_test.cpp")
FOCUS_SRC = os.path.join(ROOT, "tests", "cpp", "focus_facade_test.cpp")


def build_facade_test(out_dir, SRC=SRC):
    exe = os.path.join(str(out_dir), os.path.basename(SRC)[:-4])
    libdir = os.path.join(ROOT, "cornerstone_octree_b200")
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), SRC,
           "-L" + libdir, "-lcstone_b200", "-Wl,-rpath," + libdir, "-o", exe]
    subprocess.check_call(cmd)
    return exe


def test_facade_compiles_with_host_compiler(tmp_path):
    exe = build_facade_test(tmp_path)
    # no arguments: prints the usage and returns 2 without touching a device
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stdout


def test_focus_facade_compiles_with_host_compiler(tmp_path):
    exe = build_facade_test(tmp_path, FOCUS_SRC)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stdout


@pytest.mark.gpu
def test_focus_facade_known_answers(tmp_path):
    """groups + single-rank focus tree through focus.hpp on the reference tests' known answers"""
    exe = build_facade_test(tmp_path, FOCUS_SRC)
    r = subprocess.run([exe, "run"], capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "FOCUS_FACADE_OK" in r.stdout


@pytest.mark.gpu
def test_facade_single_rank(tmp_path):
    exe = build_facade_test(tmp_path)
    r = subprocess.run([exe, "0", "1", str(tmp_path / "id"), "4000"], capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "FACADE_OK" in r.stdout


@pytest.mark.gpu
def test_facade_two_ranks(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    exe = build_facade_test(tmp_path)
    n = 6000
    procs = [subprocess.Popen([exe, str(r), "2", str(tmp_path / "id"), str(n)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    print("\n".join(o[-2000:] for o in outs))
    assert all(p.returncode == 0 for p in procs)
    for step in range(2):
        assigned = 0
        for o in outs:
            line = [l for l in o.splitlines() if " step %d " % step in l][0].split()
            assigned += int(line[line.index("assigned") + 1])
        assert assigned == n
