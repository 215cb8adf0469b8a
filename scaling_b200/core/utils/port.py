"""Free TCP port lookup (reference: core/utils/port.py:1-16).

The reference asks the kernel for any port (``bind(0)``) and closes the socket again.  That hands out a port of the
EPHEMERAL range, which any outgoing connection (NCCL / gloo bootstrap sockets, a neighbouring job) may take before the
rendezvous store binds it — seen as sporadic EADDRINUSE when many short jobs start on one host.  Here a bindable port
below that range is preferred; ``bind(0)`` is the fallback."""
import os
import random
import socket


def _ephemeral_low() -> int:
    try:
        with open("/proc/sys/net/ipv4/ip_local_port_range") as f:
            return int(f.read().split()[0])
    except (OSError, ValueError, IndexError):
        return 32768


def find_free_port() -> int:
    high = min(_ephemeral_low(), 60000)
    low = max(10000, high - 12000)
    rng = random.Random(os.getpid() ^ int.from_bytes(os.urandom(4), "little"))
    for _ in range(200 if high > low else 0):
        port = rng.randrange(low, high)
        with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
            try:
                s.bind(("127.0.0.1", port))
            except OSError:
                continue
            return port
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        s.bind(("127.0.0.1", 0))
        return int(s.getsockname()[1])
