"""The callers either side of the index actor (SURVEY.md §8f N2 + N3), as thin host code.

  * `ItemsMonitor`  — the last-write-wins filter in front of the index (src/monitor_items.rs:56-80):
    a row reaches the index only if its write timestamp is strictly newer than the newest one seen
    for that primary key; a row without an embedding is a delete.
  * `Engine`        — the registry of index actors by id (src/engine.rs:86-123): AddIndex never
    replaces, a factory error leaves the registry untouched, DelIndex drops the actor.
  * `HttpServer`    — the vector routes the reference's integration client speaks
    (tests/integration/httpclient.rs:27-80): GET /api/v1/indexes, POST …/{keyspace}/{index}/ann,
    GET …/{keyspace}/{index}/count; 404 for an unknown index, 500 + the actor's error text for a
    failed query (src/httproutes.rs:132-150).

No vector arithmetic happens here: everything goes through an index actor (`IndexActor` /
`NativeActor`), which goes through the C ABI. Not a control plane: no database, no schema
discovery, no OpenAPI document; those stay with the reference service.
"""
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

from .index import AnnError


class DbEmbedding:
    """`struct DbEmbedding` (used at src/monitor_items.rs:57-60): primary key, optional embedding, write time."""
    __slots__ = ("primary_key", "embedding", "timestamp")

    def __init__(self, primary_key, embedding, timestamp):
        self.primary_key, self.embedding, self.timestamp = primary_key, embedding, timestamp


class ItemsMonitor:
    """`monitor_items::add` — src/monitor_items.rs:56-80."""

    def __init__(self, index):
        self.index = index
        self.timestamps = {}     # HashMap<PrimaryKey, Timestamp> (:57)
        self.dropped = 0

    def add(self, item: DbEmbedding) -> bool:
        seen = self.timestamps.get(item.primary_key)
        if seen is not None and not (seen < item.timestamp):      # :63-69 — equal timestamps do not modify
            self.dropped += 1
            return False
        self.timestamps[item.primary_key] = item.timestamp        # :65 / :70
        if item.embedding is not None:
            self.index.add_or_replace(item.primary_key, item.embedding)   # :74-75
        else:
            self.index.remove(item.primary_key)                           # :76-77
        return True

    def feed(self, items):
        for it in items:
            self.add(it)


class Engine:
    """`engine::new` message loop — src/engine.rs:86-123. `factory(index_id) -> actor` may raise."""

    def __init__(self, factory):
        self._factory = factory
        self._indexes = {}
        self._lock = threading.Lock()    # the reference serialises these through one channel

    def get_index_ids(self):             # Engine::GetIndexIds — :93-98
        with self._lock:
            return list(self._indexes.keys())

    def add_index(self, index_id):       # Engine::AddIndex — :100-115
        with self._lock:
            if index_id in self._indexes:
                return False             # "trying to replace index": ignored
            try:
                actor = self._factory(index_id)
            except Exception:
                return False             # "unable to create an index": registry unchanged
            self._indexes[index_id] = actor
            return True

    def del_index(self, index_id):       # Engine::DelIndex — :117-120
        with self._lock:
            actor = self._indexes.pop(index_id, None)
        if actor is not None and hasattr(actor, "close"):
            actor.close()                # dropping the last Sender ends the actor (usearch.rs:120,133)

    def get_index(self, index_id):       # Engine::GetIndex — :122-126
        with self._lock:
            return self._indexes.get(index_id)


def pivot_primary_keys(columns, primary_keys):
    """Vec<PrimaryKey> -> {column: [values…]} (PostIndexAnnResponse.primary_keys, tests/integration/httpclient.rs:19-20,62-65)."""
    out = {c: [] for c in columns}
    for pk in primary_keys:
        vals = pk if isinstance(pk, (tuple, list)) else (pk,)
        if len(vals) != len(columns):
            raise AnnError(f"primary key has {len(vals)} values for {len(columns)} columns")
        for c, v in zip(columns, vals):
            out[c].append(v)
    return out


class HttpServer:
    """Listens on (host, port) — port 0 picks a free one, as the integration test does (tests/integration/usearch.rs:41-46).

    `pk_columns(index_id) -> [column names]` plays DbIndex::GetPrimaryKeyColumns (src/db_index.rs:46-64).
    """

    def __init__(self, engine: Engine, pk_columns, host="127.0.0.1", port=0):
        self.engine, self.pk_columns = engine, pk_columns
        outer = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):   # quiet
                pass

            def _send(self, status, body, ctype="application/json"):
                data = body if isinstance(body, bytes) else body.encode()
                self.send_response(status)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(data)))
                self.end_headers()
                self.wfile.write(data)

            def _index(self, parts):
                return outer.engine.get_index(f"{parts[0]}.{parts[1]}")    # IndexId = "{keyspace}.{index}" (lib.rs IndexId::new)

            def do_GET(self):
                parts = [p for p in self.path.split("?")[0].split("/") if p]
                if parts == ["api", "v1", "indexes"]:
                    return self._send(200, json.dumps(outer.engine.get_index_ids()))
                if len(parts) == 6 and parts[:3] == ["api", "v1", "indexes"] and parts[5] == "count":
                    index = self._index(parts[3:5])
                    if index is None:
                        return self._send(404, "", "text/plain")
                    return self._send(200, json.dumps(int(index.count())))
                self._send(404, "", "text/plain")

            def do_POST(self):
                parts = [p for p in self.path.split("?")[0].split("/") if p]
                n = int(self.headers.get("Content-Length") or 0)
                raw = self.rfile.read(n) if n else b""
                if not (len(parts) == 6 and parts[:3] == ["api", "v1", "indexes"] and parts[5] == "ann"):
                    return self._send(404, "", "text/plain")
                index = self._index(parts[3:5])
                if index is None:                                            # httproutes.rs:137-139
                    return self._send(404, "", "text/plain")
                try:
                    req = json.loads(raw or b"{}")
                    embedding = req["embedding"]
                    limit = int(req.get("limit", 1))                         # #[serde(default)] Limit = 1 (lib.rs Limit::default)
                    if limit <= 0 or not isinstance(embedding, list):
                        raise ValueError("limit must be non-zero")
                except (ValueError, KeyError, TypeError) as e:               # axum's Json extractor rejection
                    return self._send(422, f"Failed to deserialize the JSON body into the target type: {e}", "text/plain")
                try:
                    pks, dists = index.ann(embedding, limit)
                    body = {"primary_keys": pivot_primary_keys(outer.pk_columns(f"{parts[3]}.{parts[4]}"), pks),
                            "distances": [float(d) for d in dists]}
                except AnnError as err:                                      # httproutes.rs:141-146
                    return self._send(500, f"index.ann request error: {err}", "text/plain")
                self._send(200, json.dumps(body))

        self._httpd = ThreadingHTTPServer((host, port), Handler)
        self._httpd.daemon_threads = True
        self.addr = self._httpd.server_address
        self._thread = threading.Thread(target=self._httpd.serve_forever, name="vs-http", daemon=True)
        self._thread.start()

    def close(self):
        self._httpd.shutdown()
        self._httpd.server_close()
        self._thread.join(timeout=5)
