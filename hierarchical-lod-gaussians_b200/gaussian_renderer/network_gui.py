"""Placeholder for gaussian_renderer/network_gui.py (a TCP link to a remote viewer, OUT OF SCOPE: SURVEY §2.1 row 10 / §8). It exists so
that `from gaussian_renderer import render, network_gui` of the reference's training scripts resolves; no socket is ever opened and
`conn` stays None, which is the state those scripts treat as "no viewer attached" (train_single.py:72-91)."""
host = "127.0.0.1"
port = 6009
conn = None
addr = None


def init(wish_host, wish_port):
    global host, port
    host, port = wish_host, wish_port


def try_connect():
    return None


def receive():
    raise RuntimeError("network_gui: no viewer link in this build (networking is out of scope)")


def send(message_bytes, verify):
    raise RuntimeError("network_gui: no viewer link in this build (networking is out of scope)")
