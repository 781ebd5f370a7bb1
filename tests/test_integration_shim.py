"""The reference-side binding integration/cstone_b200_shim.cpp: it compiles against the UNMODIFIED reference headers
with the host compiler, defines the symbols a client of the reference's GPU library leaves undefined, and —
on a GPU — a client translation unit of the reference (integration/shim_kat.cpp: reference headers only) reproduces
the reference's CPU results stage by stage through those names."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
SHIM = os.path.join(ROOT, "integration", "cstone_b200_shim.cpp")
KAT = os.path.join(ROOT, "integration", "_build", "shim_kat")
CUDA_INC = "/usr/local/cuda/include"

# undefined symbols a client TU of the reference really emits (probed with g++ -c + nm -u)
SURVEY_SYMBOLS = [
    "_ZN6cstone17computeSfcKeysGpuI10StrongTypeImNS_13HilbertKeyTagEEdEEvPKT0_S6_S6_PT_mRKNS_3BoxIS4_EE",
    "_ZN6cstone20computeNodeCountsGpuImEEvPKT_PjiS3_S3_jb",
    "_ZN6cstone14buildOctreeGpuImEEvPKT_NS_10OctreeViewIS1_EE",
    "_ZN6cstone12findHalosGpuImfdEEvPKT_PKiS5_S3_PKT0_RKNS_3BoxIT1_EEiiPi",
    "_ZN6cstone12sortByKeyGpuImjEEvPT_S2_PT0_S2_S4_Pvm",
]
# one name per entry point of the six reference .cu files the shim replaces
ENTRY_POINTS = ["computeSfcKeysGpu", "fillGpu", "scaleGpu", "incrementGpu", "gatherGpu", "scatterGpu", "MinMaxGpu",
                "maxNormSquareGpu", "lowerBoundGpu", "segmentMax", "reduceGpu", "sequenceGpu", "sortGpu",
                "sortByKeyTempStorage", "sortByKeyGpu", "exclusiveScanGpu", "inclusiveScanGpu", "countGpu",
                "computeNodeCountsGpu", "computeNodeOpsGpu", "rebalanceTreeGpu", "countSfcGapsGpu", "fillSfcGapsGpu",
                "buildOctreeGpu", "upsweepSumGpu", "findHalosGpu", "markMacsGpu", "gatherRanges"]


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "include", "cstone")), reason="reference tree not present")
@pytest.mark.skipif(not os.path.isdir(CUDA_INC), reason="CUDA runtime headers not present")
def test_shim_compiles_against_reference_and_defines_the_symbols(tmp_path):
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    obj = str(tmp_path / "shim.o")
    cmd = [cxx, "-std=c++20", "-DUSE_CUDA", "-Wall", "-Wno-deprecated-declarations", "-I" + os.path.join(REF, "include"),
           "-I" + os.path.join(ROOT, "include"), "-I" + CUDA_INC, "-c", SHIM, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    defined = set()
    for line in subprocess.run(["nm", obj], capture_output=True, text=True, check=True).stdout.splitlines():
        parts = line.split()
        if len(parts) == 3 and parts[1] in "TW":
            defined.add(parts[2])
    for sym in SURVEY_SYMBOLS:
        assert sym in defined, sym
    demangled = subprocess.run(["nm", "-C", obj], capture_output=True, text=True, check=True).stdout
    for name in ENTRY_POINTS:
        assert ("cstone::" + name) in demangled, name
    # nothing of the reference's own GPU library may be left undefined in the shim (it IS that library now)
    undefined = [l.split()[-1] for l in subprocess.run(["nm", "-u", obj], capture_output=True, text=True).stdout.splitlines()]
    assert not [u for u in undefined if u.startswith("_ZN6cstone")], undefined


@pytest.mark.gpu
def test_reference_client_runs_the_pipeline_through_the_shim():
    if not os.path.exists(KAT):
        pytest.skip("integration/_build/shim_kat not built (needs the reference tree at build time)")
    r = subprocess.run([KAT], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SHIM KAT OK" in r.stdout, (r.stdout[-3000:], r.stderr[-2000:])
