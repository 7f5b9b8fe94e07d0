"""Helpers for tests that drive the C++ binaries."""
import os
import socket
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# PSB_TEST_BIN=build/bin-address (or bin-thread) runs the same tests on the sanitizer binaries
BIN = os.path.abspath(os.environ.get("PSB_TEST_BIN") or os.path.join(ROOT, "build", "bin"))


def ensure_binaries():
    need = ["coordinator", "parameter_server", "worker_main", "psb_pbtool", "psb_selftest"]
    if not all(os.path.exists(os.path.join(BIN, n)) for n in need):
        subprocess.run(["make", "-C", ROOT, "-j8", "binaries"], check=True, capture_output=True)
    return BIN


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def wait_port(port, timeout=10.0):
    end = time.time() + timeout
    while time.time() < end:
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.2).close()
            return True
        except OSError:
            time.sleep(0.05)
    return False


class Proc:
    """A background binary whose output is captured to a file."""

    def __init__(self, args, cwd, log, env=None):
        self.log_path = log
        self.f = open(log, "wb")
        e = dict(os.environ)
        e.update(env or {})
        self.p = subprocess.Popen(args, cwd=cwd, stdout=self.f, stderr=subprocess.STDOUT, env=e)

    def stop(self):
        if self.p.poll() is None:
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.p.kill()
        self.f.close()

    def output(self):
        self.f.flush()
        with open(self.log_path, "r", errors="replace") as f:
            return f.read()
