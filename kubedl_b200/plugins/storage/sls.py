"""Aliyun Log Service (SLS) event backend: Kubernetes Events of the jobs shipped to an SLS logstore.

What the reference does with `--event-storage aliyun-sls` (reference
pkg/storage/backends/events/aliyun_sls/sls_logstore.go:45-279, config.go:40-74): one log per Event with the
fields FirstTimestamp / LastTimestamp / Count / Name / Kind / ObjUID / ObjNamespace / ObjName / Type / Reason /
Message (/ Region), written with PutLogs under a retry envelope (10 tries; 800 ms hold on quota errors, 200 ms on
server errors), read back with GetHistograms + GetLogs in pages of 100, sorted by FirstTimestamp and de-duplicated
by involved-object UID.  The reference links the Go SDK; here the three REST calls are spoken directly:

    POST /logstores/<store>/shards/lb          body = protobuf LogGroup            (PutLogs)
    GET  /logstores/<store>?type=histogram&…   header x-log-count                  (GetHistograms)
    GET  /logstores/<store>?type=log&…         body = JSON list of {field: value}  (GetLogs)

against `https://<project>.<endpoint>`, signed the SLS way (`Authorization: LOG <key id>:<base64 hmac-sha1>` over
verb, Content-MD5, Content-Type, Date, the sorted x-log-* / x-acs-* headers and the canonical resource).
Configuration is the reference's: SLS_ENDPOINT, SLS_KEY_ID, SLS_KEY_SECRET, SLS_PROJECT, SLS_LOG_STORE, all
required.  Without SLS_ENDPOINT the operator keeps the local JSONL log store (backends.JsonlEventBackend).
"""
from __future__ import annotations

import base64
import datetime as _dt
import hashlib
import hmac
import json
import os
import time
import urllib.error
import urllib.parse
import urllib.request
from email.utils import formatdate
from typing import Dict, List, Optional, Tuple

from . import dmo
from .backends import EventStorageBackend

API_VERSION = "0.6.0"
MAX_LINES = 100                      # SLS returns at most 100 logs per GetLogs call
RETRIES = 10
QUOTA_HOLD_S, SERVER_HOLD_S = 0.8, 0.2
QUOTA_ERRORS = ("WriteQuotaExceed", "ProjectQuotaExceed", "ShardWriteQuotaExceed")
SERVER_ERRORS = ("InternalServerError", "ServerBusy")


class SlsError(RuntimeError):
    def __init__(self, status: int, code: str, message: str):
        super().__init__(f"sls: {status} {code}: {message}")
        self.status, self.code = status, code


# ---------------------------------------------------------------------- protobuf (LogGroup, write side only)
def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _field_bytes(num: int, payload: bytes) -> bytes:      # wire type 2 (length-delimited)
    return _varint((num << 3) | 2) + _varint(len(payload)) + payload


def _field_varint(num: int, value: int) -> bytes:          # wire type 0
    return _varint(num << 3) + _varint(value)


def encode_log_group(logs: List[Tuple[int, List[Tuple[str, str]]]], topic: str = "", source: str = "") -> bytes:
    """LogGroup{ Logs=1: Log{ Time=1 (uint32), Contents=2: Content{Key=1, Value=2} }, Topic=3, Source=4 }."""
    out = bytearray()
    for ts, contents in logs:
        body = bytearray(_field_varint(1, int(ts) & 0xFFFFFFFF))
        for k, v in contents:
            body += _field_bytes(2, _field_bytes(1, k.encode()) + _field_bytes(2, v.encode()))
        out += _field_bytes(1, bytes(body))
    out += _field_bytes(3, topic.encode())
    out += _field_bytes(4, source.encode())
    return bytes(out)


def decode_log_group(buf: bytes) -> dict:
    """Inverse of encode_log_group (used by the tests' stand-in server, and handy for debugging)."""
    def fields(b: bytes):
        i = 0
        while i < len(b):
            key, i = _read_varint(b, i)
            num, wt = key >> 3, key & 7
            if wt == 0:
                v, i = _read_varint(b, i)
            elif wt == 2:
                n, i = _read_varint(b, i)
                v, i = b[i:i + n], i + n
            else:
                raise ValueError(f"unexpected wire type {wt}")
            yield num, v
    group = {"logs": [], "topic": "", "source": ""}
    for num, v in fields(buf):
        if num == 1:
            log = {"time": 0, "contents": {}}
            for n2, v2 in fields(v):
                if n2 == 1:
                    log["time"] = v2
                elif n2 == 2:
                    kv = dict(fields(v2))
                    log["contents"][kv.get(1, b"").decode()] = kv.get(2, b"").decode()
            group["logs"].append(log)
        elif num == 3:
            group["topic"] = v.decode()
        elif num == 4:
            group["source"] = v.decode()
    return group


def _read_varint(b: bytes, i: int) -> Tuple[int, int]:
    shift = n = 0
    while True:
        c = b[i]
        i += 1
        n |= (c & 0x7F) << shift
        if not c & 0x80:
            return n, i
        shift += 7


# ---------------------------------------------------------------------- signing + transport
def sign(secret: str, verb: str, content_md5: str, content_type: str, date: str, headers: Dict[str, str],
         resource: str, query: Dict[str, str]) -> str:
    canon_headers = "\n".join(f"{k}:{v}" for k, v in sorted((k.lower(), v) for k, v in headers.items()
                                                              if k.lower().startswith(("x-log-", "x-acs-"))))
    canon_resource = resource
    if query:
        canon_resource += "?" + "&".join(f"{k}={v}" for k, v in sorted(query.items()))
    msg = "\n".join([verb, content_md5, content_type, date, canon_headers, canon_resource])
    return base64.b64encode(hmac.new(secret.encode(), msg.encode(), hashlib.sha1).digest()).decode()


class SlsClient:
    def __init__(self, endpoint: str, key_id: str, key_secret: str, timeout: float = 10.0):
        ep = endpoint if "://" in endpoint else "https://" + endpoint
        u = urllib.parse.urlsplit(ep)
        self.scheme, self.host = u.scheme, u.netloc
        self.key_id, self.key_secret, self.timeout = key_id, key_secret, timeout
        # a host that is an address (tests, private gateways) cannot take a project sub-domain
        self._bare = self.host.split(":")[0].replace(".", "").isdigit() or self.host.startswith("localhost")

    def _request(self, project: str, verb: str, resource: str, query: Optional[Dict[str, str]] = None,
                 body: bytes = b"", content_type: str = "") -> Tuple[Dict[str, str], bytes]:
        query = {k: str(v) for k, v in (query or {}).items()}
        host = self.host if self._bare else f"{project}.{self.host}"
        headers = {"x-log-apiversion": API_VERSION, "x-log-signaturemethod": "hmac-sha1",
                   "x-log-bodyrawsize": str(len(body)), "Host": f"{project}.{self.host}" if self._bare else host,
                   "Date": formatdate(usegmt=True), "User-Agent": "kubedl-b200"}
        md5 = ""
        if body:
            md5 = hashlib.md5(body).hexdigest().upper()
            headers["Content-MD5"] = md5
            headers["Content-Type"] = content_type
        sig = sign(self.key_secret, verb, md5, content_type if body else "", headers["Date"], headers, resource, query)
        headers["Authorization"] = f"LOG {self.key_id}:{sig}"
        url = f"{self.scheme}://{host}{resource}"
        if query:
            url += "?" + urllib.parse.urlencode(sorted(query.items()))
        req = urllib.request.Request(url, data=body or None, method=verb, headers=headers)
        try:
            with urllib.request.urlopen(req, timeout=self.timeout) as r:
                return {k.lower(): v for k, v in r.headers.items()}, r.read()
        except urllib.error.HTTPError as e:
            raw = e.read()
            try:
                d = json.loads(raw or b"{}")
            except ValueError:
                d = {}
            raise SlsError(e.code, d.get("errorCode", "HTTPError"), d.get("errorMessage", raw.decode(errors="replace")[:200]))
        except (urllib.error.URLError, OSError) as e:
            raise SlsError(0, "RequestError", str(e))

    def put_logs(self, project: str, logstore: str, logs, topic: str = "", source: str = "") -> None:
        body = encode_log_group(logs, topic, source)
        self._request(project, "POST", f"/logstores/{logstore}/shards/lb", body=body, content_type="application/x-protobuf")

    def get_histogram_count(self, project: str, logstore: str, frm: int, to: int, query: str, topic: str = "") -> int:
        h, _ = self._request(project, "GET", f"/logstores/{logstore}",
                             {"type": "histogram", "from": frm, "to": to, "query": query, "topic": topic})
        return int(h.get("x-log-count", "0") or 0)

    def get_logs(self, project: str, logstore: str, frm: int, to: int, query: str, lines: int, offset: int,
                 reverse: bool = False, topic: str = "") -> List[Dict[str, str]]:
        _, body = self._request(project, "GET", f"/logstores/{logstore}",
                                {"type": "log", "from": frm, "to": to, "query": query, "topic": topic, "line": lines,
                                 "offset": offset, "reverse": "true" if reverse else "false"})
        out = json.loads(body or b"[]")
        return out if isinstance(out, list) else out.get("logs", [])


# ---------------------------------------------------------------------- backend
def _unix(ts: Optional[str], default: float) -> int:
    if not ts:
        return int(default)
    try:
        return int(_dt.datetime.fromisoformat(ts.replace("Z", "+00:00")).timestamp())
    except ValueError:
        return int(default)


def config_from_env(env=os.environ) -> Tuple[str, str, str, str, str]:
    """(endpoint, key id, key secret, project, logstore); every one is required (reference config.go:48-74)."""
    vals = []
    for var, what in (("SLS_ENDPOINT", "endpoint"), ("SLS_KEY_ID", "key id"), ("SLS_KEY_SECRET", "key secret"),
                      ("SLS_PROJECT", "project name"), ("SLS_LOG_STORE", "log store")):
        v = env.get(var, "")
        if not v:
            raise ValueError(f"empty sls {what} ({var})")
        vals.append(v)
    return tuple(vals)   # type: ignore[return-value]


class SlsEventBackend(EventStorageBackend):
    def __init__(self, env=os.environ, sleep=time.sleep):
        self.env, self._sleep = env, sleep
        self.client: Optional[SlsClient] = None
        self.project = self.logstore = ""

    def name(self) -> str:
        return "aliyun-sls"

    def initialize(self) -> None:
        if self.client is not None:
            return
        endpoint, key_id, key_secret, self.project, self.logstore = config_from_env(self.env)
        self.client = SlsClient(endpoint, key_id, key_secret)

    def close(self) -> None:
        self.client = None

    @staticmethod
    def wrap(event: dict, region: str):
        """One Event -> (unix time, [(field, value)...], source) in the reference's field order."""
        e = dmo.convert_event_to_dmo_event(event, region)
        contents = [("FirstTimestamp", e.first_timestamp or ""), ("LastTimestamp", e.last_timestamp or ""),
                    ("Count", str(e.count)), ("Name", e.name), ("Kind", e.kind), ("ObjUID", e.obj_uid),
                    ("ObjNamespace", e.obj_namespace), ("ObjName", e.obj_name), ("Type", e.type),
                    ("Reason", e.reason), ("Message", e.message)]
        if region:
            contents.append(("Region", region))
        src = event.get("source") or {}
        return _unix(e.last_timestamp, time.time()), contents, f"{src.get('component', '')}/{src.get('host', '')}"

    def save_event(self, event: dict, region: str) -> None:
        ts, contents, source = self.wrap(event, region)
        last: Optional[Exception] = None
        for _ in range(RETRIES):
            try:
                self.client.put_logs(self.project, self.logstore, [(ts, contents)], topic="", source=source)
                return
            except SlsError as e:
                last = e
                if e.code in QUOTA_ERRORS:
                    self._sleep(QUOTA_HOLD_S)
                elif e.code in SERVER_ERRORS:
                    self._sleep(SERVER_HOLD_S)
        raise last   # type: ignore[misc]

    def list_event(self, job_namespace: str, job_name: str, frm: Optional[str] = None, to: Optional[str] = None) -> List[dmo.Event]:
        t0, t1 = _unix(frm, 0), _unix(to, time.time() + 1)
        query = f"{job_namespace} AND {job_name}"
        total = self.client.get_histogram_count(self.project, self.logstore, t0, t1, query)
        raw: List[Dict[str, str]] = []
        offset = 0
        while offset < total:
            page = self.client.get_logs(self.project, self.logstore, t0, t1, query, min(MAX_LINES, total - offset), offset)
            if not page:
                break
            raw += page
            offset += MAX_LINES
        raw.sort(key=lambda d: d.get("FirstTimestamp", ""))
        out, seen = [], set()
        for d in raw:
            uid = d.get("ObjUID", "")
            if uid in seen:          # one entry per involved object, the earliest (reference unwrapSLSLogs)
                continue
            seen.add(uid)
            out.append(dmo.Event(name=d.get("Name", ""), kind=d.get("Kind", ""), type=d.get("Type", ""),
                                 obj_namespace=d.get("ObjNamespace", ""), obj_name=d.get("ObjName", ""), obj_uid=uid,
                                 reason=d.get("Reason", ""), message=d.get("Message", ""), count=int(d.get("Count", "0") or 0),
                                 region=d.get("Region"), first_timestamp=d.get("FirstTimestamp") or None,
                                 last_timestamp=d.get("LastTimestamp") or None))
        return out
