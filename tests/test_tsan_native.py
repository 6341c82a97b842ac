"""ThreadSanitizer run of the C++ parameter-server transport, bundle writer and queues.
The only suppressed reports are the lock-free optimizer applies / value copies that implement the
reference's Hogwild semantics (csrc/native/tsan.supp)."""
import os
import subprocess

import pytest

from distributed_tensorflow_b200.utils import build

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parameter_server_is_race_free_under_thread_sanitizer(tmp_path):
    try:
        exe = build.build_tsan_stress(tmp_path / "dtf_tsan_stress")
    except Exception as e:                                      # toolchain without libtsan
        pytest.skip("cannot build with -fsanitize=thread: %s" % e)
    env = dict(os.environ, TSAN_OPTIONS="suppressions=%s halt_on_error=1 exitcode=66" % os.path.join(REPO, "csrc/native/tsan.supp"))
    p = subprocess.run([str(exe), "60", str(tmp_path)], env=env, capture_output=True, text=True, timeout=300)
    assert "ThreadSanitizer" not in p.stderr, p.stderr[-3000:]
    assert p.returncode == 0, (p.returncode, p.stderr[-2000:])
    assert "tsan_stress ok: 360 steps from 6 threads" in p.stdout
