"""The host-side hand-off protocol between the trackers' threads (csrc/seq_flags.h, used by csrc/sequence.cu) under
ThreadSanitizer: tests/cpp/flags_tsan_test.cpp replays the protocol of seq_work on plain memory, so a missing happens-before
edge is a reported data race, a wrong waiter is a lost wake-up (time-out) and a wrong order is a counted violation.
(compute-sanitizer is not available on the GPU pool; this covers the part of the hand-off that lives on the host.)"""
import os
import shutil
import subprocess

import pytest

from conftest import ROOT


@pytest.mark.parametrize("trackers", [2, 7, 20])
def test_handoff_flags_under_thread_sanitizer(tmp_path, trackers):
    cxx = shutil.which("g++")
    if not cxx:
        pytest.skip("no g++")
    exe = str(tmp_path / "flags_tsan_test")
    build = subprocess.run([cxx, "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-pthread", "-I", os.path.join(ROOT, "densemonoslam_b200", "csrc"),
                            os.path.join(ROOT, "tests", "cpp", "flags_tsan_test.cpp"), "-o", exe], capture_output=True, text=True)
    if build.returncode != 0:
        pytest.skip("ThreadSanitizer build not available: " + build.stderr[-300:])
    run = subprocess.run([exe, str(trackers), "400", "3"], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0, run.stdout[-500:] + run.stderr[-2000:]
    assert "WARNING: ThreadSanitizer" not in run.stderr, run.stderr[-3000:]
    assert "ordering violations=0" in run.stdout
