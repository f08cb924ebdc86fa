"""CPU test of the BVH8 builder + the host build of the traversal core (the same __host__ __device__ code the
CUDA kernels use) against brute force.  Compiles tests/cpp/bvh8_check.cu with nvcc as a host program."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bvh8_builder_and_traversal_match_bruteforce():
    exe = os.path.join(ROOT, "build", "bvh8_check")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    src = [os.path.join(ROOT, "tests", "cpp", "bvh8_check.cu"),
           os.path.join(ROOT, "software-path-tracer_b200", "csrc", "bvh_build.cpp")]
    subprocess.run(["nvcc", "-O2", "-std=c++17", "-fmad=false", "-Xcompiler", "-ffp-contract=off", "-diag-suppress", "20014",
                    "-Wno-deprecated-gpu-targets", "-o", exe] + src + ["-lpthread"], check=True, capture_output=True)
    out = subprocess.run([exe, "1500"], capture_output=True, text=True)
    print(out.stdout)
    assert out.returncode == 0 and out.stdout.strip().endswith("OK")
