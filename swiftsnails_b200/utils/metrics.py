"""Process-local metrics: counters, gauges and rate meters with a Prometheus text exposition.

The reference has no metrics at all — glog banners only (SURVEY.md §5.5;
/root/reference/src/core/system/master/terminate.h:95-97, worker/terminate.h:30-32).  Here every
long-running object (trainer, table, framework roles) reports into one registry.  Everything is
host-side arithmetic on values the host already has (tokens per step, keys per request, the loss
that was copied back anyway), so recording a sample never synchronises with the GPU.

    from swiftsnails_b200.utils.metrics import registry
    registry().counter("w2v_tokens_total").inc(n)
    print(registry().render())                 # Prometheus text format
    registry().serve(9100)                     # optional: HTTP endpoint (stdlib, daemon thread)
"""
from __future__ import annotations

import collections
import threading
import time


class Counter:
    __slots__ = ("name", "help", "_v", "_lock")

    def __init__(self, name, help=""):
        self.name, self.help, self._v, self._lock = name, help, 0.0, threading.Lock()

    def inc(self, v=1.0):
        if v < 0:
            raise ValueError("a counter only goes up")
        with self._lock:
            self._v += v

    @property
    def value(self):
        return self._v

    kind = "counter"


class Gauge:
    __slots__ = ("name", "help", "_v")

    def __init__(self, name, help=""):
        self.name, self.help, self._v = name, help, 0.0

    def set(self, v):
        self._v = float(v)

    @property
    def value(self):
        return self._v

    kind = "gauge"


class Meter:
    """Events per second over a sliding window (default 5 s) — e.g. words/s of a trainer."""

    __slots__ = ("name", "help", "window", "_samples", "_lock", "_total")
    kind = "gauge"

    def __init__(self, name, help="", window=5.0):
        self.name, self.help, self.window = name, help, float(window)
        self._samples, self._lock, self._total = collections.deque(), threading.Lock(), 0.0

    def mark(self, n=1.0, now=None):
        now = time.monotonic() if now is None else now
        with self._lock:
            self._total += n
            self._samples.append((now, n))
            lo = now - self.window
            while self._samples and self._samples[0][0] < lo:
                self._samples.popleft()

    def rate(self, now=None):
        now = time.monotonic() if now is None else now
        with self._lock:
            s = [(t, n) for t, n in self._samples if t >= now - self.window]
            if len(s) < 2:
                return 0.0
            span = s[-1][0] - s[0][0]
            return sum(n for _, n in s[1:]) / span if span > 0 else 0.0

    @property
    def value(self):
        return self.rate()


class Registry:
    def __init__(self):
        self._m, self._lock, self._server = {}, threading.Lock(), None

    def _get(self, cls, name, help, **kw):
        with self._lock:
            m = self._m.get(name)
            if m is None:
                m = self._m[name] = cls(name, help, **kw)
            elif not isinstance(m, cls):
                raise TypeError("metric %r already registered as %s" % (name, type(m).__name__))
            return m

    def counter(self, name, help="") -> Counter:
        return self._get(Counter, name, help)

    def gauge(self, name, help="") -> Gauge:
        return self._get(Gauge, name, help)

    def meter(self, name, help="", window=5.0) -> Meter:
        return self._get(Meter, name, help, window=window)

    def snapshot(self):
        with self._lock:
            return {k: m.value for k, m in sorted(self._m.items())}

    def render(self):
        """Prometheus text exposition format (version 0.0.4)."""
        out = []
        with self._lock:
            items = sorted(self._m.items())
        for name, m in items:
            if m.help:
                out.append("# HELP %s %s" % (name, m.help))
            out.append("# TYPE %s %s" % (name, m.kind))
            out.append("%s %.17g" % (name, m.value))
        return "\n".join(out) + "\n"

    def serve(self, port, addr="127.0.0.1"):
        """Serve ``render()`` at http://addr:port/metrics from a daemon thread; returns the bound port."""
        import http.server

        reg = self

        class H(http.server.BaseHTTPRequestHandler):
            def do_GET(self):
                body = reg.render().encode()
                self.send_response(200)
                self.send_header("Content-Type", "text/plain; version=0.0.4")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def log_message(self, *a):
                pass

        srv = http.server.ThreadingHTTPServer((addr, port), H)
        threading.Thread(target=srv.serve_forever, daemon=True).start()
        self._server = srv
        return srv.server_address[1]

    def stop(self):
        if self._server is not None:
            self._server.shutdown()
            self._server.server_close()
            self._server = None

    def clear(self):
        with self._lock:
            self._m.clear()


_REGISTRY = Registry()


def registry() -> Registry:
    return _REGISTRY
