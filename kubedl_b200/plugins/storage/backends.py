"""Storage backends: interfaces, query, registry and two implementations.

Interfaces: reference pkg/storage/backends/interface.go:31-72 (ObjectStorageBackend,
EventStorageBackend), query.go:25-41; registry: backends/registry/registry.go:26-116.

  "mysql"        object backend.  Same tables/columns as the reference's gorm models
                 (job_info, replica_info; reference backends/objects/mysql/mysql.go:46-443,
                 env MYSQL_HOST/PORT/DB_NAME/USER/PASSWORD, config.go:41-62).  With MYSQL_HOST set and
                 a DB-API driver installed (pymysql / MySQLdb / mysql.connector) it connects to that
                 server; neither exists on this box, where the same statements (plain ANSI SQL, `?`
                 placeholders rewritten for the driver) run on sqlite3 (file or :memory:,
                 `MYSQL_DB_NAME` = path).  The reference's
                 queries against non-existent columns (gmt_create / gmt_submit / is_del,
                 mysql.go:113,192-195,212) are written against the real ones.
  "aliyun-sls"   event backend.  The reference ships events to an Aliyun SLS logstore with
                 <=10 retries (sls_logstore.go:45-279, env SLS_ENDPOINT/KEY_ID/KEY_SECRET/PROJECT/
                 LOG_STORE); there is no network here, so the logstore is a local JSONL file
                 (`SLS_LOG_STORE` = path) with the same retry envelope around the write; with SLS_ENDPOINT
                 set, `sls.SlsEventBackend` ships them to that logstore over the SLS REST API instead.
"""
from __future__ import annotations

import abc
import json
import os
import sqlite3
import threading
import time
from dataclasses import asdict, dataclass, fields
from typing import Dict, List, Optional

from ...api import common as c
from . import dmo


@dataclass
class QueryPagination:
    page_num: int = 1
    page_size: int = 20
    count: int = 0


@dataclass
class Query:
    job_id: str = ""
    name: str = ""
    namespace: str = ""
    region: str = ""
    status: str = ""
    start_time: Optional[str] = None
    end_time: Optional[str] = None
    is_del: Optional[int] = None
    pagination: Optional[QueryPagination] = None


class ObjectStorageBackend(abc.ABC):
    @abc.abstractmethod
    def initialize(self) -> None: ...
    @abc.abstractmethod
    def close(self) -> None: ...
    @abc.abstractmethod
    def name(self) -> str: ...
    @abc.abstractmethod
    def save_pod(self, pod: dict, default_container_name: str, region: str) -> None: ...
    @abc.abstractmethod
    def list_pods(self, job_id: str, region: str) -> List[dmo.Pod]: ...
    @abc.abstractmethod
    def stop_pod(self, ns: str, name: str, pod_id: str) -> None: ...
    @abc.abstractmethod
    def save_job(self, job: dict, kind: str, specs: Dict[str, dict], job_status: dict, region: str) -> None: ...
    @abc.abstractmethod
    def get_job(self, ns: str, name: str, job_id: str, region: str) -> Optional[dmo.Job]: ...
    @abc.abstractmethod
    def list_jobs(self, query: Query) -> List[dmo.Job]: ...
    @abc.abstractmethod
    def stop_job(self, ns: str, name: str, job_id: str, region: str) -> None: ...
    @abc.abstractmethod
    def delete_job(self, ns: str, name: str, job_id: str, region: str) -> None: ...


class EventStorageBackend(abc.ABC):
    @abc.abstractmethod
    def initialize(self) -> None: ...
    @abc.abstractmethod
    def close(self) -> None: ...
    @abc.abstractmethod
    def name(self) -> str: ...
    @abc.abstractmethod
    def save_event(self, event: dict, region: str) -> None: ...
    @abc.abstractmethod
    def list_event(self, job_namespace: str, job_name: str, frm: Optional[str], to: Optional[str]) -> List[dmo.Event]: ...


def _cols(cls) -> List[str]:
    return [f.name for f in fields(cls)]


class _Db:
    """The little of DB-API the backend needs, over sqlite3 or a MySQL driver: `execute` with `?`
    placeholders (rewritten to the driver's `%s`), `commit`, `close`."""

    def __init__(self, conn, paramstyle: str = "qmark", dialect: str = "sqlite"):
        self.conn, self.paramstyle, self.dialect = conn, paramstyle, dialect

    def execute(self, sql: str, args=()):
        if self.paramstyle in ("format", "pyformat"):
            sql = sql.replace("%", "%%").replace("?", "%s")
        cur = self.conn.cursor()
        cur.execute(sql, tuple(args))
        return cur

    def commit(self) -> None:
        self.conn.commit()

    def close(self) -> None:
        self.conn.close()


def mysql_driver():
    """The first importable MySQL DB-API module (none ships in this image), else None."""
    import importlib
    for name in ("pymysql", "MySQLdb", "mysql.connector"):
        try:
            return importlib.import_module(name)
        except ImportError:
            continue
    return None


def open_object_db(path: Optional[str] = None, env=os.environ, driver=None) -> _Db:
    """MYSQL_HOST set and a driver present -> that MySQL server (MYSQL_HOST / MYSQL_PORT / MYSQL_DB_NAME /
    MYSQL_USER / MYSQL_PASSWORD, reference backends/objects/mysql/config.go:41-62); otherwise sqlite3
    on `path` / MYSQL_DB_NAME / memory."""
    host = env.get("MYSQL_HOST", "")
    driver = driver if driver is not None else (mysql_driver() if host else None)
    if host and driver is not None and not path:
        kw = dict(host=host, port=int(env.get("MYSQL_PORT", "3306") or 3306), user=env.get("MYSQL_USER", "root"),
                  password=env.get("MYSQL_PASSWORD", ""), database=env.get("MYSQL_DB_NAME", "kubedl"))
        return _Db(driver.connect(**kw), getattr(driver, "paramstyle", "format"), "mysql")
    return _Db(sqlite3.connect(path or env.get("MYSQL_DB_NAME") or ":memory:", check_same_thread=False), "qmark", "sqlite")


class SqlObjectBackend(ObjectStorageBackend):
    """job_info / replica_info with upsert guarded by resourceVersion (newer wins) and terminal
    status stickiness, like reference mysql.go:291-414."""

    def __init__(self, path: Optional[str] = None, driver=None):
        self.path = path
        self._driver = driver
        self._db: Optional[_Db] = None
        self._lock = threading.RLock()
        self._initialized = False

    def name(self) -> str:
        return "mysql"

    def initialize(self) -> None:
        with self._lock:
            if self._initialized:
                return
            self._db = open_object_db(self.path, driver=self._driver)
            mysql = self._db.dialect == "mysql"
            keyed = ("namespace", "name", "job_id", "pod_id")            # indexed: bounded type on MySQL
            for cls in (dmo.Job, dmo.Pod):
                def coltype(n):
                    if n in ("deleted", "is_in_etcd"):
                        return "INTEGER"
                    return "VARCHAR(255)" if mysql and n in keyed else "TEXT"
                cols = ", ".join(f"{n} {coltype(n)}" for n in _cols(cls))
                pk = "id BIGINT PRIMARY KEY AUTO_INCREMENT" if mysql else "id INTEGER PRIMARY KEY AUTOINCREMENT"
                self._db.execute(f"CREATE TABLE IF NOT EXISTS {cls.TABLE} ({pk}, {cols})")
            for ddl in ("CREATE INDEX {ine}idx_job ON job_info (namespace, name, job_id)",
                        "CREATE INDEX {ine}idx_pod ON replica_info (namespace, name, pod_id)"):
                try:
                    self._db.execute(ddl.format(ine="" if mysql else "IF NOT EXISTS "))
                except Exception:          # MySQL has no IF NOT EXISTS for indexes: "duplicate key name" on restart
                    if not mysql:
                        raise
            self._db.commit()
            self._initialized = True

    def close(self) -> None:
        with self._lock:
            if self._db is not None:
                self._db.close()
                self._db = None
                self._initialized = False

    # ---- helpers
    def _row(self, cls, where: str, args) -> Optional[object]:
        cur = self._db.execute(f"SELECT {', '.join(_cols(cls))} FROM {cls.TABLE} WHERE {where}", args)
        r = cur.fetchone()
        return cls(*r) if r else None

    def _rows(self, cls, where: str, args, tail: str = "") -> list:
        cur = self._db.execute(f"SELECT {', '.join(_cols(cls))} FROM {cls.TABLE} WHERE {where} {tail}", args)
        return [cls(*r) for r in cur.fetchall()]

    def _insert(self, obj) -> None:
        d = asdict(obj)
        d["gmt_modified"] = c.now_rfc3339()
        cols = list(d)
        self._db.execute(f"INSERT INTO {obj.TABLE} ({', '.join(cols)}) VALUES ({', '.join('?' * len(cols))})",
                         [d[k] for k in cols])
        self._db.commit()

    def _update(self, obj, where: str, args) -> None:
        d = asdict(obj)
        d["gmt_modified"] = c.now_rfc3339()
        sets = ", ".join(f"{k} = ?" for k in d)
        self._db.execute(f"UPDATE {obj.TABLE} SET {sets} WHERE {where}", list(d.values()) + list(args))
        self._db.commit()

    @staticmethod
    def _newer(old_version: str, new_version: str) -> bool:
        try:
            return int(new_version) >= int(old_version)
        except (TypeError, ValueError):
            return True

    # ---- pods
    def save_pod(self, pod: dict, default_container_name: str, region: str) -> None:
        new = dmo.convert_pod_to_dmo_pod(pod, default_container_name, region)
        with self._lock:
            old = self._row(dmo.Pod, "namespace = ? AND name = ? AND pod_id = ?", (new.namespace, new.name, new.pod_id))
            if old is None:
                self._insert(new)
                return
            if not self._newer(old.version, new.version):
                return
            if old.status in (c.POD_SUCCEEDED, c.POD_FAILED):
                new.status, new.gmt_finished, new.remark = old.status, old.gmt_finished or new.gmt_finished, old.remark or new.remark
            if old.gmt_started and not new.gmt_started:
                new.gmt_started = old.gmt_started
            self._update(new, "namespace = ? AND name = ? AND pod_id = ?", (new.namespace, new.name, new.pod_id))

    def list_pods(self, job_id: str, region: str) -> List[dmo.Pod]:
        with self._lock:
            where, args = "job_id = ?", [job_id]
            if region:
                where += " AND deploy_region = ?"
                args.append(region)
            return self._rows(dmo.Pod, where, args, "ORDER BY replica_type, gmt_created DESC, name")

    def stop_pod(self, ns: str, name: str, pod_id: str) -> None:
        with self._lock:
            old = self._row(dmo.Pod, "namespace = ? AND name = ? AND pod_id = ?", (ns, name, pod_id))
            if old is None:
                return
            if old.status not in (c.POD_SUCCEEDED, c.POD_FAILED):
                old.status = "Stopped"
                old.gmt_finished = old.gmt_finished or c.now_rfc3339()
            old.is_in_etcd = 0
            self._update(old, "namespace = ? AND name = ? AND pod_id = ?", (ns, name, pod_id))

    # ---- jobs
    def save_job(self, job: dict, kind: str, specs: Dict[str, dict], job_status: dict, region: str) -> None:
        new = dmo.convert_job_to_dmo_job(job, kind, specs, job_status, region)
        with self._lock:
            old = self._row(dmo.Job, "namespace = ? AND name = ? AND job_id = ?", (new.namespace, new.name, new.job_id))
            if old is None:
                self._insert(new)
                return
            if not self._newer(old.version, new.version):
                return
            if old.status in (c.JOB_SUCCEEDED, c.JOB_FAILED):
                new.status, new.gmt_finished = old.status, old.gmt_finished or new.gmt_finished
            self._update(new, "namespace = ? AND name = ? AND job_id = ?", (new.namespace, new.name, new.job_id))

    def get_job(self, ns: str, name: str, job_id: str, region: str) -> Optional[dmo.Job]:
        with self._lock:
            where, args = "namespace = ? AND name = ? AND job_id = ?", [ns, name, job_id]
            if region:
                where += " AND deploy_region = ?"
                args.append(region)
            return self._row(dmo.Job, where, args)

    def list_jobs(self, query: Query) -> List[dmo.Job]:
        where, args = ["1 = 1"], []
        for col, val in (("job_id", query.job_id), ("name", query.name), ("namespace", query.namespace),
                         ("deploy_region", query.region), ("status", query.status)):
            if val:
                where.append(f"{col} = ?")
                args.append(val)
        if query.start_time:
            where.append("gmt_created >= ?")
            args.append(query.start_time)
        if query.end_time:
            where.append("gmt_created <= ?")
            args.append(query.end_time)
        if query.is_del is not None:
            where.append("deleted = ?")
            args.append(query.is_del)
        with self._lock:
            w = " AND ".join(where)
            if query.pagination is not None:
                query.pagination.count = self._db.execute(f"SELECT COUNT(*) FROM job_info WHERE {w}", args).fetchone()[0]
                tail = f"ORDER BY gmt_created DESC LIMIT {int(query.pagination.page_size)} " \
                       f"OFFSET {int((query.pagination.page_num - 1) * query.pagination.page_size)}"
            else:
                tail = "ORDER BY gmt_created DESC"
            return self._rows(dmo.Job, w, args, tail)

    def stop_job(self, ns: str, name: str, job_id: str, region: str) -> None:
        with self._lock:
            old = self.get_job(ns, name, job_id, region)
            if old is None:
                return
            if old.status not in (c.JOB_SUCCEEDED, c.JOB_FAILED):
                old.status = "Stopped"
                old.gmt_finished = old.gmt_finished or c.now_rfc3339()
            old.is_in_etcd = 0
            self._update(old, "namespace = ? AND name = ? AND job_id = ?", (ns, name, job_id))

    def delete_job(self, ns: str, name: str, job_id: str, region: str) -> None:
        with self._lock:
            old = self.get_job(ns, name, job_id, region)
            if old is None:
                return
            old.deleted, old.is_in_etcd = 1, 0
            self._update(old, "namespace = ? AND name = ? AND job_id = ?", (ns, name, job_id))


class JsonlEventBackend(EventStorageBackend):
    MAX_RETRIES = 10

    def __init__(self, path: Optional[str] = None):
        self.path = path or os.environ.get("SLS_LOG_STORE") or "/tmp/kubedl-b200/events.jsonl"
        self._lock = threading.Lock()
        self._fh = None

    def name(self) -> str:
        return "aliyun-sls"

    def initialize(self) -> None:
        os.makedirs(os.path.dirname(self.path) or ".", exist_ok=True)
        self._fh = open(self.path, "a", buffering=1)

    def close(self) -> None:
        if self._fh:
            self._fh.close()
            self._fh = None

    def save_event(self, event: dict, region: str) -> None:
        d = asdict(dmo.convert_event_to_dmo_event(event, region))
        line = json.dumps(d, sort_keys=True)
        last = None
        for attempt in range(self.MAX_RETRIES):
            try:
                with self._lock:
                    self._fh.write(line + "\n")
                return
            except OSError as e:   # "server error": short hold, then retry
                last = e
                time.sleep(0.2)
        raise last

    def list_event(self, job_namespace, job_name, frm=None, to=None) -> List[dmo.Event]:
        out = []
        try:
            with open(self.path) as f:
                for line in f:
                    d = json.loads(line)
                    if d["obj_namespace"] != job_namespace:
                        continue
                    if not (d["obj_name"] == job_name or d["obj_name"].startswith(job_name + "-")):
                        continue
                    ts = d.get("last_timestamp") or ""
                    if (frm and ts < frm) or (to and ts > to):
                        continue
                    out.append(dmo.Event(**d))
        except FileNotFoundError:
            pass
        return out


# ---------------------------------------------------------------------- registry
class _Registry:
    def __init__(self):
        self._lock = threading.Lock()
        self.objects: Dict[str, ObjectStorageBackend] = {}
        self.events: Dict[str, EventStorageBackend] = {}

    def add_object_backend(self, b: ObjectStorageBackend) -> None:
        with self._lock:
            self.objects[b.name()] = b

    def get_object_backend(self, name: str) -> Optional[ObjectStorageBackend]:
        with self._lock:
            return self.objects.get(name)

    def remove_object_backend(self, name: str) -> None:
        with self._lock:
            self.objects.pop(name, None)

    def add_event_backend(self, b: EventStorageBackend) -> None:
        with self._lock:
            self.events[b.name()] = b

    def get_event_backend(self, name: str) -> Optional[EventStorageBackend]:
        with self._lock:
            return self.events.get(name)

    def remove_event_backend(self, name: str) -> None:
        with self._lock:
            self.events.pop(name, None)


REGISTRY = _Registry()


def register_storage_backends() -> None:
    """reference registry.go:32-43 (called from main.go:97)."""
    if REGISTRY.get_object_backend("mysql") is None:
        REGISTRY.add_object_backend(SqlObjectBackend())
    if REGISTRY.get_event_backend("aliyun-sls") is None:
        if os.environ.get("SLS_ENDPOINT"):
            from .sls import SlsEventBackend          # a real logstore is configured: speak to it
            REGISTRY.add_event_backend(SlsEventBackend())
        else:
            REGISTRY.add_event_backend(JsonlEventBackend())
