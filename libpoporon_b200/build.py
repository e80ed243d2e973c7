"""In-tree build of the native libraries (nvcc for sm_100a + gcc), driven by the Makefiles in csrc/."""
from __future__ import annotations

import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_dir() -> str:
    return os.path.join(_HERE, "lib")


def build_native(verbose: bool = False, jobs: int = 8) -> None:
    """Compile libpoporon_cuda.so (``-gencode arch=compute_100a,code=sm_100a -lineinfo``) and
    libpoporon.so.  nvcc cross-compiles without a GPU, so this also runs on the CPU-only build box."""
    out = None if verbose else subprocess.DEVNULL
    env = dict(os.environ)
    env.setdefault("PATH", "")
    if "/usr/local/cuda/bin" not in env["PATH"]:
        env["PATH"] = "/usr/local/cuda/bin:" + env["PATH"]
    subprocess.run(["make", "-C", os.path.join(_HERE, "csrc", "cuda"), f"-j{jobs}"], check=True, stdout=out, env=env)
    subprocess.run(["make", "-C", os.path.join(_HERE, "csrc", "host")], check=True, stdout=out, env=env)


if __name__ == "__main__":
    build_native(verbose=True)
