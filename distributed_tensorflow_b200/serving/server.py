"""JSON-over-HTTP predict endpoint for exported models (README.md:29-41 of the reference).

The reference serves its SavedModel with the external ``simple_tensorflow_serving`` tool::

    simple_tensorflow_serving --port=8500 --model_base_path="./saved_model"
    curl -H "Content-Type: application/json" -X POST \
         -d '{"keys": [[11.0], [2.0]], "features": [[1], [2]]}' http://127.0.0.1:8500

This module is that server: it picks the highest-numbered version directory under
``--model_base_path``, reads the ``serving_default`` signature, feeds the request's top-level keys
(or the ``{"data": {...}}`` envelope) as signature inputs -- casting to the declared dtypes, so the
float ``keys`` of the README become int32 -- and answers with the signature outputs as nested lists:
``{"keys": [[11], [2]], "prediction": [[w+b], [2w+b]]}``.

Run: ``python -m distributed_tensorflow_b200.serving --port=8500 --model_base_path=./saved_model``
"""
from __future__ import annotations

import json
import os
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, Optional

import numpy as np

from ..export.saved_model import DEFAULT_SERVING_SIGNATURE_DEF_KEY, load_saved_model

_DT_TO_NP = {1: np.float32, 3: np.int32, 9: np.int64}


def latest_version_dir(model_base_path: str) -> str:
    versions = [d for d in os.listdir(model_base_path)
                if d.isdigit() and os.path.isdir(os.path.join(model_base_path, d))]
    if not versions:
        raise FileNotFoundError("no numeric version directory under %r" % model_base_path)
    return os.path.join(model_base_path, max(versions, key=int))


class ModelServer:
    """Loaded model + predict(); independent of the HTTP layer so it can be benchmarked directly."""

    def __init__(self, model_base_path: str, signature: str = DEFAULT_SERVING_SIGNATURE_DEF_KEY,
                 device: str = "auto"):
        self.model_base_path = model_base_path
        self.export_dir = latest_version_dir(model_base_path)
        self.version = int(os.path.basename(self.export_dir))
        loaded = load_saved_model(self.export_dir)
        if signature not in loaded["signatures"]:
            raise KeyError("signature %r not in %s" % (signature, sorted(loaded["signatures"])))
        self.signature = loaded["signatures"][signature]
        self.variables: Dict[str, np.ndarray] = loaded["variables"]
        self.description = loaded["description"] or {}
        self.input_dtypes = {k: _DT_TO_NP[v.dtype] for k, v in self.signature.inputs.items()}
        self.output_names = sorted(self.signature.outputs)
        self.model_name = self.description.get("model", "linear")
        # wide-and-deep: the sparse ids ride in `features` (or in an extra integer input `ids`); keep
        # float64 up to the predictor so that ids >= 2**24 are not rounded by a float32 cast
        self.optional_inputs: Dict[str, type] = {}
        if self.model_name == "wide_deep":
            self.input_dtypes["features"] = np.float64
            self.optional_inputs["ids"] = np.int64
        self._impl = self._build(device)
        self.requests = 0

    def _build(self, device: str):
        if self.model_name == "linear":
            from ..models.linear import LinearModel
            m = LinearModel()
            return lambda inputs: m.predict(self.variables, inputs)
        if self.model_name == "wide_deep":
            from ..models.wide_deep_serving import WideDeepPredictor
            p = WideDeepPredictor(self.variables, self.description.get("config", {}), device=device)
            return p.predict
        raise ValueError("exported model %r is not servable by this build" % self.model_name)

    def predict(self, request: dict) -> dict:
        data = request.get("data", request) if isinstance(request, dict) else None
        if not isinstance(data, dict):
            raise ValueError("request body must be a JSON object")
        inputs = {}
        for alias, dt in self.input_dtypes.items():
            if alias not in data:
                raise ValueError("missing input %r (signature inputs: %s)" % (alias, sorted(self.input_dtypes)))
            inputs[alias] = np.asarray(data[alias]).astype(dt)
        for alias, dt in self.optional_inputs.items():
            if alias in data:
                inputs[alias] = np.asarray(data[alias]).astype(dt)
        out = self._run(inputs)
        return {k: np.asarray(out[k]).tolist() for k in self.output_names}

    def _run(self, inputs: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
        n = {v.shape[0] for v in inputs.values() if v.ndim > 0}
        if len(n) > 1:
            raise ValueError("inputs disagree on batch size: %s" % sorted(n))
        out = self._impl(inputs)
        self.requests += 1
        return out

    def predict_json(self, body: bytes) -> bytes:
        """Request body -> response body.  Bodies made of rectangular numeric arrays (the normal case) are
        parsed and formatted by the C++ fast path (``csrc/native/json_fast.cc``); anything else goes through
        the generic ``json`` module with identical results."""
        fast = _FastJson.get()
        if fast is not None and body and b'"data"' not in body:
            inputs = {}
            for alias, dt in self.input_dtypes.items():
                arr = fast.parse(body, alias)
                if arr is None:
                    inputs = None
                    break
                inputs[alias] = arr.astype(dt)
            if inputs is not None:
                for alias, dt in self.optional_inputs.items():
                    if ('"%s"' % alias).encode() in body:
                        arr = fast.parse(body, alias)
                        if arr is None:
                            inputs = None
                            break
                        inputs[alias] = arr.astype(dt)
            if inputs is not None:
                out = self._run(inputs)
                parts = []
                for k in self.output_names:
                    txt = fast.format(np.asarray(out[k]))
                    if txt is None:
                        parts = None
                        break
                    parts.append(json.dumps(k).encode() + b": " + txt)
                if parts is not None:
                    return b"{" + b", ".join(parts) + b"}"
                return json.dumps({k: np.asarray(out[k]).tolist() for k in self.output_names}).encode()
        return json.dumps(self.predict(json.loads(body or b"{}"))).encode()


class _FastJson:
    """ctypes binding of the native JSON fast path; ``get()`` returns None when the library is missing."""

    _inst = None
    _tried = False

    @classmethod
    def get(cls):
        if not cls._tried:
            cls._tried = True
            try:
                from .. import _native
                cls._inst = cls(_native.lib())
            except Exception:       # noqa: BLE001  (no native library: generic path only)
                cls._inst = None
        return cls._inst

    def __init__(self, L):
        import ctypes as C
        self.C, self.L = C, L
        L.dtfn_json_numeric_array.restype = C.c_longlong
        L.dtfn_json_numeric_array.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_void_p, C.c_size_t,
                                              C.POINTER(C.c_longlong), C.POINTER(C.c_int)]
        L.dtfn_json_format_array.restype = C.c_longlong
        L.dtfn_json_format_array.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_longlong), C.c_int, C.c_void_p, C.c_size_t]

    def parse(self, body: bytes, key: str):
        C = self.C
        cap = len(body) // 2 + 1                     # every number takes at least two bytes ("1,")
        buf = np.empty(cap, np.float64)
        dims = (C.c_longlong * 4)()
        nd = C.c_int(0)
        n = self.L.dtfn_json_numeric_array(body, len(body), key.encode(), buf.ctypes.data, cap, dims, C.byref(nd))
        if n < 0:
            return None
        shape = tuple(int(dims[i]) for i in range(nd.value))
        if int(np.prod(shape)) != n:
            return None
        return buf[:n].reshape(shape)

    def format(self, arr: np.ndarray):
        C = self.C
        if arr.ndim < 1 or arr.ndim > 4:
            return None
        if np.issubdtype(arr.dtype, np.integer):
            data, kind = np.ascontiguousarray(arr, np.int64), 1
        elif np.issubdtype(arr.dtype, np.floating):
            data, kind = np.ascontiguousarray(arr, np.float32), 0
        else:
            return None
        dims = (C.c_longlong * 4)(*arr.shape)
        cap = data.size * (26 + 2 * arr.ndim) + 16
        out = C.create_string_buffer(cap)
        n = self.L.dtfn_json_format_array(data.ctypes.data, kind, dims, arr.ndim, out, cap)
        if n <= 0:
            return None
        return out.raw[:n]


def make_http_server(model: ModelServer, host: str = "0.0.0.0", port: int = 8500) -> ThreadingHTTPServer:
    class Handler(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"
        # Response head and body must leave in ONE segment on keep-alive connections: with an unbuffered wfile
        # the body is a second small write that Nagle holds back until the client's delayed ACK (40 ms) -- measured
        # 44 ms per request for any body under ~1 MB.  Buffered writes (flushed once per request) + TCP_NODELAY.
        wbufsize = 1 << 20

        def setup(self):
            super().setup()
            try:
                import socket as _socket
                self.connection.setsockopt(_socket.IPPROTO_TCP, _socket.TCP_NODELAY, 1)
            except OSError:
                pass

        def _send(self, code: int, payload: dict) -> None:
            self._send_raw(code, json.dumps(payload).encode())

        def _send_raw(self, code: int, body: bytes) -> None:
            self.send_response(code)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def do_GET(self):  # noqa: N802  (health / model info)
            self._send(200, {"model": model.model_name, "version": model.version,
                             "signature_inputs": sorted(model.input_dtypes), "signature_outputs": model.output_names,
                             "requests": model.requests})

        def do_POST(self):  # noqa: N802
            try:
                n = int(self.headers.get("Content-Length", "0"))
                self._send_raw(200, model.predict_json(self.rfile.read(n)))
            except (ValueError, KeyError, TypeError) as e:
                self._send(400, {"error": str(e)})
            except Exception as e:  # pragma: no cover
                self._send(500, {"error": "%s: %s" % (type(e).__name__, e)})

        def log_message(self, fmt, *args):
            pass

    return ThreadingHTTPServer((host, port), Handler)


def main(argv=None) -> int:
    import argparse
    ap = argparse.ArgumentParser(description="JSON predict server for exported models")
    ap.add_argument("--port", type=int, default=8500)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--model_base_path", default="./saved_model")
    ap.add_argument("--device", default="auto")
    args = ap.parse_args(argv)
    ms = ModelServer(args.model_base_path, device=args.device)
    httpd = make_http_server(ms, args.host, args.port)
    print("Serving model %s version %d from %s on port %d" % (ms.model_name, ms.version, ms.export_dir, args.port),
          flush=True)
    try:
        httpd.serve_forever()
    except KeyboardInterrupt:
        pass
    finally:
        httpd.server_close()
    return 0
