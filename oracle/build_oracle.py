"""Builds the C oracle (oracle/ksim_oracle.c) into oracle/_build/libks_oracle.so with gcc.
TEST INFRASTRUCTURE: building the checker is not using it; the product never loads this library."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_build", "libks_oracle.so")


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "ksim_oracle.c")
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(src):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    # -ffp-contract=off: CPython never fuses a*b+c, and the comparison is bit for bit
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-o", OUT, src, "-lm"], check=True)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
