"""Minimal ``POST /parse`` front end over the C ABI, for drop-in trials against the operator.

Same verb, path and media types as the reference's client interface
(LogParserRestClient.java:23-25, :37-39); the operator is pointed here through
``quarkus.rest-client.log-parser.url`` (operator-deployment.yaml:50-51).  The production host is
the Java resource sketched in INTEGRATION.md; this module only proves the contract end to end.

    python -m operator_b200.service --patterns /shared/patterns --port 8080
"""
from __future__ import annotations

import argparse
import json
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

from .engine import Engine, ScanError
from . import patterns


def make_handler(engine: Engine):
    class Handler(BaseHTTPRequestHandler):
        def do_POST(self):
            if self.path.rstrip("/") != "/parse":
                self.send_error(404)
                return
            body = self.rfile.read(int(self.headers.get("Content-Length", "0")))
            try:
                out = json.dumps(engine.parse(body)).encode("utf-8")
                code = 200
            except ScanError as e:      # the operator sees a failed Uni (PodFailureWatcher.java:295-304)
                out = json.dumps({"error": str(e)}).encode("utf-8")
                code = 400 if e.status == -2 else 500
            self.send_response(code)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(out)))
            self.end_headers()
            self.wfile.write(out)

        def log_message(self, *a):
            pass

    return Handler


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--patterns", default=patterns.PATTERN_CACHE_DIR)
    ap.add_argument("--port", type=int, default=8080)
    args = ap.parse_args()
    engine = Engine()
    # the library the operator's sync step compiled, if it is there and current; else compile now and leave it
    if patterns.load_compiled(engine, args.patterns) is None:
        patterns.validate_patterns(engine, args.patterns, save=True)
    ThreadingHTTPServer(("0.0.0.0", args.port), make_handler(engine)).serve_forever()


if __name__ == "__main__":
    main()
