"""Generates tests/golden/sph_l3.json from the vendored sphericart CPU code (build container only:
needs /root/reference).  Usage: python oracle/gen_golden.py"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
S = "/root/reference/submodules/sphericart/sphericart"


def main():
    os.makedirs(os.path.join(HERE, "_ref"), exist_ok=True)
    exe = os.path.join(HERE, "_ref", "sph_ref")
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", f"-I{S}/include",
                           os.path.join(HERE, "sph_ref_main.cpp"), f"{S}/src/sphericart.cpp", "-o", exe])
    out = subprocess.check_output([exe])
    os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
    with open(os.path.join(ROOT, "tests", "golden", "sph_l3.json"), "wb") as f:
        f.write(out)
    print("wrote tests/golden/sph_l3.json", len(out), "bytes")


if __name__ == "__main__":
    main()
