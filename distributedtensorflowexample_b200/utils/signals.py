"""Signal handling of the reference (main.py:50-55): SIGHUP/SIGINT/SIGTERM raise
``SystemExit(128 + signum)``, which unwinds through ``managed_session`` on workers (stopping the
Supervisor services) and breaks the ps's sleep loop."""
from __future__ import annotations

import signal
import sys

HANDLED = (signal.SIGHUP, signal.SIGINT, signal.SIGTERM)


def _shutdown(signum, frame):
    sys.exit(128 + signum)


def install_shutdown_handlers() -> None:
    for s in HANDLED:
        signal.signal(s, _shutdown)
