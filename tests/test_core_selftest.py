"""Native concurrency self-test of the C++ control plane (actor mailboxes under concurrent tell/ask/stop, timers, the
reader/writer lease table, the journal store). Runs as an ordinary binary everywhere, and under ThreadSanitizer where
the toolchain ships libtsan (the reference runs no race detector at all)."""
import pathlib
import subprocess

import pytest

from determined_b200 import _build

ROOT = pathlib.Path(__file__).resolve().parent.parent


def test_selftest_binary(tmp_path):
    _build.build_core()
    out = subprocess.run([str(ROOT / "determined_b200" / "_lib" / "det-selftest"), str(tmp_path / "store")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "selftest ok" in out.stdout, out.stdout + out.stderr


def test_selftest_under_thread_sanitizer(tmp_path):
    try:
        exe = _build.build_selftest_tsan()
    except RuntimeError as e:
        if "tsan" in str(e).lower():
            pytest.skip("libtsan is not installed in this image")
        raise
    out = subprocess.run([str(exe), str(tmp_path / "store")], capture_output=True, text=True, timeout=900,
                         env={"TSAN_OPTIONS": "halt_on_error=0 exitcode=66"})
    assert "WARNING: ThreadSanitizer" not in out.stderr, out.stderr[-4000:]
    assert out.returncode == 0, out.stdout + out.stderr[-2000:]
