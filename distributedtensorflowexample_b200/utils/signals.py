"""Signal handling of the reference (main.py:50-55): SIGHUP/SIGINT/SIGTERM raise
``SystemExit(128 + signum)``, which unwinds through ``managed_session`` on workers (stopping the
Supervisor services) and breaks the ps's sleep loop.

Python runs a signal handler only when the main thread is back in the interpreter.  A task that is stuck
inside native code - a rendezvous waiting for peers that will never come, an exchange during which the
shutdown signals are masked - would therefore ignore SIGTERM until that call returns (minutes).  A watcher
thread closes the gap: the C-level handler reports every signal through a wake-up descriptor at once; if the
process is still alive ``grace_s`` seconds after a shutdown signal, the watcher ends it with the exit code
the handler would have produced.
"""
from __future__ import annotations

import os
import signal
import socket
import sys
import threading
import time

HANDLED = (signal.SIGHUP, signal.SIGINT, signal.SIGTERM)
_watcher = None


def _shutdown(signum, frame):
    sys.exit(128 + signum)


def _watch(rsock: socket.socket, grace_s: float) -> None:
    handled = {int(s) for s in HANDLED}
    while True:
        try:
            data = rsock.recv(64)
        except OSError:
            return
        if not data:
            return
        hit = [b for b in data if b in handled]
        if hit:
            time.sleep(grace_s)                      # the normal path (SystemExit in the main thread) wins if it can
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(128 + hit[0])


def install_shutdown_handlers(grace_s: float = 30.0) -> None:
    """Install the reference's handlers; ``grace_s`` <= 0 disables the stuck-in-native-code watchdog."""
    global _watcher
    for s in HANDLED:
        signal.signal(s, _shutdown)
    if grace_s > 0 and _watcher is None and threading.current_thread() is threading.main_thread():
        rsock, wsock = socket.socketpair()
        wsock.setblocking(False)
        signal.set_wakeup_fd(wsock.fileno(), warn_on_full_buffer=False)
        t = threading.Thread(target=_watch, args=(rsock, grace_s), name="signal-watchdog", daemon=True)
        t._dtfe_socks = (rsock, wsock)               # keep both ends alive as long as the thread
        t.start()
        _watcher = t
