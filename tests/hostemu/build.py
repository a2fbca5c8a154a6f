"""TEST INFRASTRUCTURE ONLY: builds tests/hostemu/_build/libhostemu.so (CPU build of the kernel bodies)."""
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
SO = HERE / "_build" / "libhostemu.so"


def build():
    srcs = [HERE / "hostemu.cu"] + list((ROOT / "pulselib_b200" / "csrc").glob("*.cuh")) + [ROOT / "include" / "pulse_envs.h"]
    if SO.exists() and all(s.stat().st_mtime < SO.stat().st_mtime for s in srcs):
        return SO
    SO.parent.mkdir(exist_ok=True)
    cmd = ["nvcc", "-x", "cu", "-O2", "-std=c++17", "--fmad=false", "-shared", "-Xcompiler", "-fPIC,-ffp-contract=off",
           "-gencode", "arch=compute_100a,code=sm_100a",
           "-I", str(ROOT / "include"), "-I", str(ROOT / "pulselib_b200" / "csrc"), str(HERE / "hostemu.cu"), "-o", str(SO)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(r.stdout + r.stderr)
    return SO


if __name__ == "__main__":
    print(build())
