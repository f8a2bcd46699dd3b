"""utils/metrics.py: counters, gauges, meters, Prometheus text and the HTTP endpoint (CPU only)."""
import urllib.request

import pytest

from swiftsnails_b200.utils.metrics import Registry


def test_counter_gauge_meter_render():
    r = Registry()
    c = r.counter("tokens_total", "tokens seen")
    c.inc(5)
    c.inc(7)
    assert r.counter("tokens_total") is c and c.value == 12
    with pytest.raises(ValueError):
        c.inc(-1)
    with pytest.raises(TypeError):
        r.gauge("tokens_total")
    r.gauge("loss").set(0.25)
    m = r.meter("rate", window=10.0)
    for k in range(5):
        m.mark(100, now=1000.0 + k)          # 100 events per second
    assert abs(m.rate(now=1004.0) - 100.0) < 1e-9
    assert m.rate(now=2000.0) == 0.0         # window empty
    text = r.render()
    assert "# TYPE tokens_total counter" in text and "tokens_total 12" in text
    assert "# HELP tokens_total tokens seen" in text and "loss 0.25" in text
    assert r.snapshot()["loss"] == 0.25


def test_http_endpoint():
    r = Registry()
    r.counter("hits_total").inc(3)
    port = r.serve(0)
    try:
        body = urllib.request.urlopen("http://127.0.0.1:%d/metrics" % port, timeout=5).read().decode()
    finally:
        r.stop()
    assert "hits_total 3" in body
