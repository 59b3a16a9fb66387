"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU checkers for the sxmc hot path:
  * ``oracle.port``  -- ctypes face of oracle/sxmc_oracle.c (this repo's plain-C restatement)
  * ``oracle.ref``   -- ctypes face of oracle/_ref/libsxmc_ref.so (the unmodified reference
                        sources compiled for the CPU; only present where they were built)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package.  Nothing under sxmc_b200/ does.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = "/root/reference"


def build(verbose=False):
    """Compile the C restatement and, where the reference tree is present, oracle/_ref."""
    targets = ["port"]
    if os.path.isdir(os.path.join(REFERENCE_ROOT, "src")):
        targets.append("ref")
    out = subprocess.run(["make", "-C", HERE] + targets, capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


def have_port():
    return os.path.exists(os.path.join(HERE, "liboracle_port.so"))


def have_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "libsxmc_ref.so"))
