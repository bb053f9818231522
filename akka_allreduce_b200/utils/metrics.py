"""Prometheus text endpoint for a master or a worker.

The reference loads Akka's cluster-metrics extension (application.conf:27-33, Sigar native library in `native/`) and
never reads it; its only observable is the MBytes/sec line the worker prints.  Here the same numbers —
round, live workers, rounds completed, throughput — can be scraped: ``cluster.metricsPort`` in the config file for the
master, ``--metrics-port`` for a worker.  No dependency: the exposition format is a few lines of text.
"""
from __future__ import annotations

import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Mapping


def render(metrics: Mapping[str, float], labels: Mapping[str, str] | None = None) -> str:
    """``{'aar_round': 12}`` -> Prometheus text exposition (gauges)."""
    lab = ""
    if labels:
        lab = "{" + ",".join(f'{k}="{str(v)}"' for k, v in sorted(labels.items())) + "}"
    lines = []
    for name, value in metrics.items():
        lines.append(f"# TYPE {name} gauge")
        lines.append(f"{name}{lab} {float(value):g}")
    return "\n".join(lines) + "\n"


class MetricsExporter:
    def __init__(self, collect: Callable[[], Mapping[str, float]], port: int = 0, host: str = "127.0.0.1",
                 labels: Mapping[str, str] | None = None):
        exporter = self

        class Handler(BaseHTTPRequestHandler):
            def do_GET(self):  # noqa: N802 (http.server's naming)
                if self.path.split("?")[0] not in ("/metrics", "/"):
                    self.send_error(404)
                    return
                try:
                    body = render(exporter._collect(), exporter._labels).encode()
                except Exception as e:  # a scrape must not take the node down
                    self.send_error(500, str(e)[:100])
                    return
                self.send_response(200)
                self.send_header("Content-Type", "text/plain; version=0.0.4")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def log_message(self, *args):  # scrapes are not log lines
                pass

        self._collect, self._labels = collect, dict(labels or {})
        self._server = ThreadingHTTPServer((host, port), Handler)
        self._server.daemon_threads = True
        self.port = self._server.server_address[1]
        self._thread = threading.Thread(target=self._server.serve_forever, name="aar-metrics", daemon=True)
        self._thread.start()

    def close(self) -> None:
        self._server.shutdown()
        self._server.server_close()
        self._thread.join(timeout=5)
