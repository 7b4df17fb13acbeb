"""The FAST traversal structure on the host: every quantised node is decoded as the kernel decodes it and must contain
what lies below it with the margin the culling argument needs (csrc/fast_bvh.h).  No GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "ray-tracing-one-weekend_b200", "csrc")


@pytest.fixture(scope="module")
def checker(pkg, tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("fastbvh") / "fast_bvh_check")
    cmd = ["g++", "-O2", "-std=c++17", "-I/usr/local/cuda/include", os.path.join(ROOT, "tests", "cpp", "fast_bvh_check.cpp"),
           os.path.join(CSRC, "fast_bvh.cpp"), "-o", exe, "-L" + CSRC, "-lrtow_b200", "-Wl,-rpath," + CSRC,
           "-Wl,-rpath-link,/usr/local/cuda/lib64"]
    out = subprocess.run(cmd, capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    return exe


@pytest.mark.parametrize("scene,n", [("random_ned", 0), ("random", 0), ("bvh_stress", 1), ("bvh_stress", 7), ("bvh_stress", 20000)])
def test_decoded_nodes_contain_their_content(checker, scene, n):
    out = subprocess.run([checker, scene, str(n)], capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 0 and out.stdout.startswith("ok"), out.stdout + out.stderr
    nodes, leaves = map(int, out.stdout.split()[1:3])
    assert leaves >= max(n, 1) and nodes >= 1


def test_the_threaded_build_gives_the_same_structure_for_any_thread_count(checker):
    """build_fast_bvh builds subtrees of a big scene on the host's threads (csrc/fast_bvh.cpp); how the work is cut depends on
    the object count alone, so the node array is the same byte for byte with 1, 3 or all threads."""
    hashes = set()
    for threads in ("1", "3", ""):
        env = dict(os.environ)
        env.pop("RTOW_BVH_THREADS", None)
        if threads:
            env["RTOW_BVH_THREADS"] = threads
        out = subprocess.run([checker, "bvh_stress", "70000"], capture_output=True, text=True, cwd=ROOT, env=env)
        assert out.returncode == 0 and out.stdout.startswith("ok"), out.stdout + out.stderr
        hashes.add(out.stdout.split()[3])
    assert len(hashes) == 1
