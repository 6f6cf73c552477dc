"""CPU coverage of the multi-rank plumbing (csrc/shm_group.h): real processes meeting in the shared-memory
segment — barrier, mailboxes, and the time-out when a peer never arrives (no hang)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = tmp_path_factory.mktemp("shm") / "shm_group_check"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", os.path.join(ROOT, "tests", "shm_group_check.cpp"), "-o", str(out), "-lrt"])
    return str(out)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_allgather_rounds(exe, world):
    p = subprocess.run([exe, str(world), "2000"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    assert f"ok world={world}" in p.stdout


def test_absent_peer_times_out(exe):
    env = dict(os.environ, GE_SHM_TIMEOUT="1")
    p = subprocess.run([exe, "3", "10", "absent"], capture_output=True, text=True, timeout=60, env=env)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "timed out as expected" in p.stdout
