"""DMO converters, storage backends and persist controllers.
Case names follow reference pkg/storage/dmo/converters/{pod,job,event}_test.go."""
import json
import os

import pytest

from cp_helpers import Cluster, tmpl
from kubedl_b200.api import common as c
from kubedl_b200.api import kinds as k
from kubedl_b200.controllers import persist
from kubedl_b200.plugins.storage import backends as B
from kubedl_b200.plugins.storage import dmo

TS = "2020-03-01T10:00:00Z"


def _pod(phase=None, cstate=None, owner=True, rtype=True, containers=None, init=None, uid="pod-uid"):
    m = {"name": "job-worker-0", "namespace": "ns", "uid": uid, "resourceVersion": "3", "creationTimestamp": TS,
         "labels": {}}
    if rtype:
        m["labels"][c.REPLICA_TYPE_LABEL] = "worker"
    if owner:
        m["ownerReferences"] = [{"kind": "TFJob", "name": "job", "uid": "job-uid", "controller": True}]
    pod = {"kind": "Pod", "metadata": m,
           "spec": {"containers": containers or [{"name": "tensorflow", "image": "tf:1"}], "initContainers": init or []},
           "status": {}}
    if phase:
        pod["status"] = {"phase": phase, "podIP": "10.0.0.1", "hostIP": "192.168.0.1",
                         "containerStatuses": [{"name": "tensorflow", "state": cstate or {}}]}
    return pod


def test_pod_owner_reference_error_and_replica_type_error():
    with pytest.raises(dmo.NoDependentOwner):
        dmo.convert_pod_to_dmo_pod(_pod(owner=False), "tensorflow")
    with pytest.raises(dmo.NoReplicaTypeLabel):
        dmo.convert_pod_to_dmo_pod(_pod(rtype=False), "tensorflow")


def test_pod_status_unknown_pending_running():
    d = dmo.convert_pod_to_dmo_pod(_pod(), "tensorflow", region="cn")
    assert d.status == "Unknown" and d.job_id == "job-uid" and d.replica_type == "worker"
    assert d.deploy_region == "cn" and d.deleted == 0 and d.is_in_etcd == 1 and d.image == "tf:1"
    d = dmo.convert_pod_to_dmo_pod(_pod("Pending"), "tensorflow")
    assert d.status == "Pending" and d.gmt_started is None and d.pod_ip == "10.0.0.1" and d.host_ip == "192.168.0.1"
    d = dmo.convert_pod_to_dmo_pod(_pod("Running", {"running": {"startedAt": "2020-03-01T10:00:05Z"}}), "tensorflow")
    assert d.status == "Running" and d.gmt_started == "2020-03-01T10:00:05Z"
    d = dmo.convert_pod_to_dmo_pod(_pod("Running", {}), "tensorflow")
    assert d.gmt_started == TS                                   # falls back to creation time


def test_pod_status_succeeded_failed_with_remark():
    term = {"startedAt": "2020-03-01T10:00:05Z", "finishedAt": "2020-03-01T10:10:05Z", "exitCode": 137,
            "reason": "OOMKilled", "message": "killed"}
    d = dmo.convert_pod_to_dmo_pod(_pod("Succeeded", {"terminated": term}), "tensorflow")
    assert (d.status, d.gmt_started, d.gmt_finished, d.remark) == \
        ("Succeeded", "2020-03-01T10:00:05Z", "2020-03-01T10:10:05Z", None)
    d = dmo.convert_pod_to_dmo_pod(_pod("Failed", {"terminated": term}), "tensorflow")
    assert d.remark == "Reason: OOMKilled\nExitCode: 137\nMessage: killed"


def test_pod_single_and_multiple_container_resources():
    d = dmo.convert_pod_to_dmo_pod(_pod(containers=[{"name": "tensorflow", "image": "i", "resources": {
        "requests": {"cpu": "1", "memory": "1Gi"}, "limits": {"cpu": "2", "memory": "2Gi"}}}]), "tensorflow")
    assert json.loads(d.resources) == {"requests": {"cpu": "1", "memory": "1Gi"}, "limits": {"cpu": "2", "memory": "2Gi"}}
    d = dmo.convert_pod_to_dmo_pod(_pod(
        containers=[{"name": "a", "image": "a", "resources": {"requests": {"cpu": "1"}}},
                    {"name": "tensorflow", "image": "main", "resources": {"requests": {"cpu": "500m", "memory": "1Gi"}}}],
        init=[{"name": "init", "resources": {"requests": {"cpu": "3", "memory": "100Mi"}}}]), "tensorflow")
    assert json.loads(d.resources) == {"requests": {"cpu": "3", "memory": "1Gi"}}   # max(init, sum containers)
    assert d.image == "main"                                                          # default container's image


def _job(kind="TFJob", conds=(), annotations=None, completion=None):
    info = k.SCHEME.info(kind)
    rt = "Worker" if "Worker" in info.replica_types else info.replica_types[-1]
    job = k.new_job(kind, "j", "ns", {rt: {"replicas": 2, "template": tmpl(
        info.default_container_name, resources={"requests": {"cpu": "1"}})}})
    job["metadata"].update({"uid": "job-uid", "resourceVersion": "5", "creationTimestamp": TS})
    if annotations:
        job["metadata"]["annotations"] = annotations
    job["status"] = {"conditions": [{"type": t, "status": "True"} for t in conds], "replicaStatuses": {}}
    if completion:
        job["status"]["completionTime"] = completion
    return job


def test_job_created_status_and_resources():
    job = _job()
    kind, specs, status = dmo.extract_typed_job_infos(job)
    d = dmo.convert_job_to_dmo_job(job, kind, specs, status, "")
    assert (d.status, d.kind, d.job_id, d.version, d.tenant, d.owner, d.deploy_region) == \
        ("Created", "TFJob", "job-uid", "5", "", "", None)
    assert json.loads(d.resources) == {"Worker": {"replicas": 2, "resources": {"requests": {"cpu": "1"}}}}


def test_job_tenancy_region_annotation_and_succeeded_kinds():
    ann = {c.ANNOTATION_TENANCY_INFO: json.dumps({"tenant": "t1", "user": "u1", "region": "r1"})}
    job = _job(annotations=ann)
    d = dmo.convert_job_to_dmo_job(job, *dmo.extract_typed_job_infos(job)[:2], job["status"], "")
    assert (d.tenant, d.owner, d.deploy_region) == ("t1", "u1", "r1")
    d = dmo.convert_job_to_dmo_job(job, "TFJob", {}, job["status"], "explicit")
    assert d.deploy_region == "explicit"
    for kind in ("XDLJob", "PyTorchJob", "XGBoostJob"):
        j = _job(kind, conds=("Created", "Running", "Succeeded"), completion="2020-03-01T11:00:00Z")
        kk, specs, status = dmo.extract_typed_job_infos(j)
        d = dmo.convert_job_to_dmo_job(j, kk, specs, status, "")
        assert (d.kind, d.status, d.gmt_finished) == (kind, "Succeeded", "2020-03-01T11:00:00Z")
    with pytest.raises(ValueError):
        dmo.extract_typed_job_infos({"kind": "Deployment", "metadata": {"name": "x"}})


def test_event_conversion_keeps_involved_object():
    ev = {"metadata": {"name": "e1", "namespace": "ns"}, "type": "Normal", "reason": "SuccessfulCreatePod",
          "message": "Created pod: p", "count": 2, "firstTimestamp": TS, "lastTimestamp": TS,
          "involvedObject": {"kind": "TFJob", "namespace": "ns", "name": "j", "uid": "job-uid"}}
    d = dmo.convert_event_to_dmo_event(ev, "cn")
    assert (d.kind, d.obj_name, d.obj_uid, d.count, d.region, d.reason) == \
        ("TFJob", "j", "job-uid", 2, "cn", "SuccessfulCreatePod")


def test_sql_backend_upsert_stop_delete_and_query(tmp_path):
    b = B.SqlObjectBackend(str(tmp_path / "kubedl.db"))
    b.initialize()
    job = _job(conds=("Created",))
    b.save_job(job, "TFJob", job["spec"]["tfReplicaSpecs"], job["status"], "cn")
    assert b.get_job("ns", "j", "job-uid", "cn").status == "Created"
    job["metadata"]["resourceVersion"] = "9"
    job["status"]["conditions"].append({"type": "Succeeded"})
    job["status"]["completionTime"] = "2020-03-01T11:00:00Z"
    b.save_job(job, "TFJob", job["spec"]["tfReplicaSpecs"], job["status"], "cn")
    job["metadata"]["resourceVersion"] = "4"                               # stale write is ignored
    job["status"]["conditions"] = [{"type": "Running"}]
    b.save_job(job, "TFJob", job["spec"]["tfReplicaSpecs"], job["status"], "cn")
    got = b.get_job("ns", "j", "job-uid", "cn")
    assert (got.status, got.version, got.gmt_finished) == ("Succeeded", "9", "2020-03-01T11:00:00Z")
    pg = B.QueryPagination(page_num=1, page_size=10)
    assert [j.name for j in b.list_jobs(B.Query(namespace="ns", status="Succeeded", pagination=pg))] == ["j"]
    assert pg.count == 1 and b.list_jobs(B.Query(status="Failed")) == []
    pod = _pod("Running", {"running": {"startedAt": TS}})
    b.save_pod(pod, "tensorflow", "cn")
    assert [p.name for p in b.list_pods("job-uid", "cn")] == ["job-worker-0"]
    b.stop_pod("ns", "job-worker-0", "pod-uid")
    p = b.list_pods("job-uid", "")[0]
    assert p.status == "Stopped" and p.is_in_etcd == 0 and p.gmt_finished
    b.stop_job("ns", "j", "job-uid", "cn")                                  # terminal status is sticky
    b.delete_job("ns", "j", "job-uid", "cn")
    got = b.get_job("ns", "j", "job-uid", "cn")
    assert got.status == "Succeeded" and got.deleted == 1 and got.is_in_etcd == 0
    assert b.list_jobs(B.Query(is_del=0)) == []
    b.close()


def test_registry():
    B.register_storage_backends()
    assert B.REGISTRY.get_object_backend("mysql").name() == "mysql"
    assert B.REGISTRY.get_event_backend("aliyun-sls").name() == "aliyun-sls"
    assert B.REGISTRY.get_object_backend("nope") is None


def test_persist_controllers_full_loop(tmp_path):
    cl = Cluster()
    sql = B.SqlObjectBackend(":memory:")
    ev = B.JsonlEventBackend(str(tmp_path / "events.jsonl"))
    B.REGISTRY.add_object_backend(sql)
    B.REGISTRY.add_event_backend(ev)
    try:
        ctrls = persist.setup_with_manager(cl.mgr, "mysql", "aliyun-sls", region="b200-box", workloads="*")
        assert set(ctrls) == {"job/PyTorchJob", "job/TFJob", "job/XGBoostJob", "job/XDLJob", "pod", "event"}
        job = cl.submit(k.new_job("TFJob", "tf", "ns", {"Worker": {"replicas": 2, "restartPolicy": "Never",
                                                                    "template": tmpl("tensorflow")}}))
        uid = job["metadata"]["uid"]
        cl.kubelet.set_all("ns", "Running")
        cl.settle()
        row = sql.get_job("ns", "tf", uid, "b200-box")
        assert row.status == "Running" and row.kind == "TFJob" and row.is_in_etcd == 1
        pods = sql.list_pods(uid, "b200-box")
        assert sorted(p.name for p in pods) == ["tf-worker-0", "tf-worker-1"] and all(p.status == "Running" for p in pods)
        cl.kubelet.set_phase("ns", "tf-worker-0", "Succeeded")
        cl.kubelet.set_phase("ns", "tf-worker-1", "Succeeded")
        cl.settle()
        assert sql.get_job("ns", "tf", uid, "b200-box").status == "Succeeded"
        reasons = {e.reason for e in ev.list_event("ns", "tf")}
        assert {"SuccessfulCreatePod", "SuccessfulCreateService", "JobSucceeded"} <= reasons
        cl.api.delete("TFJob", "ns", "tf")
        cl.settle()
        row = sql.get_job("ns", "tf", uid, "b200-box")
        assert row.deleted == 1 and row.is_in_etcd == 0 and row.status == "Succeeded"
        assert all(p.is_in_etcd == 0 for p in sql.list_pods(uid, ""))
        with pytest.raises(ValueError):
            persist.setup_with_manager(cl.mgr, object_storage="no-such-backend")
        assert persist.parse_id_name(persist.gen_id_name("u", "n")) == ("u", "n")
    finally:
        B.REGISTRY.remove_object_backend("mysql")
        B.REGISTRY.remove_event_backend("aliyun-sls")


def test_mysql_driver_path_with_a_fake_driver(monkeypatch):
    """MYSQL_HOST + a DB-API driver -> the backend talks to that server.  No MySQL here, so the
    "driver" is a shim with MySQL's `%s` parameter style and DDL dialect over sqlite: it checks the
    placeholder rewriting, the DDL branch and the connection parameters."""
    import re
    import sqlite3
    from kubedl_b200.plugins.storage import backends as B
    seen = {}

    class Cursor:
        def __init__(self, cur):
            self.cur = cur

        def execute(self, sql, args=()):
            assert "?" not in sql, sql                                   # placeholders were rewritten
            seen.setdefault("sql", []).append(sql)
            sql = sql.replace("%s", "?").replace("%%", "%")
            sql = sql.replace("BIGINT PRIMARY KEY AUTO_INCREMENT", "INTEGER PRIMARY KEY AUTOINCREMENT")
            sql = re.sub(r"VARCHAR\\(\\d+\\)", "TEXT", sql)
            return self.cur.execute(sql, args)

        def fetchone(self):
            return self.cur.fetchone()

        def fetchall(self):
            return self.cur.fetchall()

    class Conn:
        def __init__(self):
            self.db = sqlite3.connect(":memory:", check_same_thread=False)

        def cursor(self):
            return Cursor(self.db.cursor())

        def commit(self):
            self.db.commit()

        def close(self):
            self.db.close()

    class Driver:
        paramstyle = "format"

        @staticmethod
        def connect(**kw):
            seen["connect"] = kw
            return Conn()

    monkeypatch.setenv("MYSQL_HOST", "db.internal")
    monkeypatch.setenv("MYSQL_PORT", "3307")
    monkeypatch.setenv("MYSQL_USER", "kubedl")
    monkeypatch.setenv("MYSQL_PASSWORD", "pw")
    monkeypatch.setenv("MYSQL_DB_NAME", "kubedl_prod")
    be = B.SqlObjectBackend(driver=Driver)
    be.initialize()
    assert seen["connect"] == {"host": "db.internal", "port": 3307, "user": "kubedl", "password": "pw", "database": "kubedl_prod"}
    assert any("AUTO_INCREMENT" in q and "VARCHAR(255)" in q for q in seen["sql"])
    be.initialize()                                                       # idempotent
    be.save_pod(_pod(phase="Running"), "tensorflow", "cn")
    be.save_pod(_pod(phase="Succeeded"), "tensorflow", "cn")
    (row,) = be.list_pods("job-uid", "cn")
    assert row.status == "Succeeded"
    be.stop_pod(row.namespace, row.name, row.pod_id)
    assert be.list_pods("job-uid", "cn")[0].is_in_etcd == 0
    be.close()
    # no driver, or no MYSQL_HOST: sqlite, as before
    monkeypatch.delenv("MYSQL_HOST")
    monkeypatch.setenv("MYSQL_DB_NAME", ":memory:")
    assert B.open_object_db().dialect == "sqlite"


# ---------------------------------------------------------------- Aliyun SLS event backend over HTTP
class _FakeSls:
    """A stand-in for the three SLS REST calls the backend makes: checks the request signature, decodes the
    protobuf LogGroup, can inject errors, answers histogram / log queries from what it stored."""

    def __init__(self, key_id="kid", secret="s3cret"):
        import http.server
        import threading
        from kubedl_b200.plugins.storage import sls
        self.key_id, self.secret, self.logs, self.fail, self.requests = key_id, secret, [], [], []
        outer = self

        class H(http.server.BaseHTTPRequestHandler):
            def log_message(self, *a):
                pass

            def _check(self, body):
                import hashlib
                import urllib.parse
                u = urllib.parse.urlsplit(self.path)
                q = dict(urllib.parse.parse_qsl(u.query, keep_blank_values=True))
                hd = {k.lower(): v for k, v in self.headers.items()}       # header names are case-insensitive
                md5 = hashlib.md5(body).hexdigest().upper() if body else ""
                assert hd.get("content-md5", "") == md5
                want = sls.sign(outer.secret, self.command, md5, hd.get("content-type", "") if body else "", hd["date"], hd, u.path, q)
                assert hd["authorization"] == f"LOG {outer.key_id}:{want}", "bad signature"
                assert hd["x-log-apiversion"] == "0.6.0" and hd["x-log-bodyrawsize"] == str(len(body))
                assert hd["host"].startswith("proj.")
                return u.path, q

            def _reply(self, code, body=b"", headers=None):
                self.send_response(code)
                for k, v in (headers or {}).items():
                    self.send_header(k, v)
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def do_POST(self):
                body = self.rfile.read(int(self.headers.get("Content-Length", 0)))
                path, _ = self._check(body)
                outer.requests.append(("POST", path))
                if outer.fail:
                    code = outer.fail.pop(0)
                    return self._reply(500 if code in sls.SERVER_ERRORS else 403,
                                       json.dumps({"errorCode": code, "errorMessage": "injected"}).encode())
                assert path == "/logstores/store/shards/lb"
                g = sls.decode_log_group(body)
                for lg in g["logs"]:
                    outer.logs.append(dict(lg["contents"], __time__=str(lg["time"]), __source__=g["source"]))
                self._reply(200)

            def do_GET(self):
                path, q = self._check(b"")
                outer.requests.append(("GET", q.get("type")))
                assert path == "/logstores/store"
                ns, _, name = q["query"].partition(" AND ")
                hit = [l for l in outer.logs if int(q["from"]) <= int(l["__time__"]) <= int(q["to"])
                       and l["ObjNamespace"] == ns and name in l["ObjName"]]
                if q["type"] == "histogram":
                    return self._reply(200, b"[]", {"x-log-count": str(len(hit)), "x-log-progress": "Complete"})
                page = hit[int(q["offset"]):int(q["offset"]) + int(q["line"])]
                assert int(q["line"]) <= 100
                self._reply(200, json.dumps(page).encode(), {"x-log-count": str(len(page)), "x-log-progress": "Complete"})

        self.srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), H)
        self.url = f"http://127.0.0.1:{self.srv.server_address[1]}"
        threading.Thread(target=self.srv.serve_forever, daemon=True).start()

    def stop(self):
        self.srv.shutdown()


def _k8s_event(i, uid, name, ns="ns1", reason="SuccessfulCreatePod"):
    ts = f"2026-01-01T00:{i // 60:02d}:{i % 60:02d}Z"
    return {"kind": "Event", "metadata": {"name": f"ev-{i}", "namespace": ns}, "type": "Normal", "reason": reason,
            "message": f"Created pod: {name}", "count": 1, "firstTimestamp": ts, "lastTimestamp": ts,
            "involvedObject": {"kind": "Pod", "namespace": ns, "name": name, "uid": uid},
            "source": {"component": "pytorch-controller", "host": "node-0"}}


def test_sls_log_group_protobuf_roundtrip():
    from kubedl_b200.plugins.storage import sls
    buf = sls.encode_log_group([(1767225600, [("Name", "ev-1"), ("Message", "héllo wörld " * 30)]),
                                (5, [("Count", "3")])], topic="", source="ctl/host")
    g = sls.decode_log_group(buf)
    assert g["source"] == "ctl/host" and g["topic"] == "" and len(g["logs"]) == 2
    assert g["logs"][0]["time"] == 1767225600 and g["logs"][0]["contents"]["Message"].startswith("héllo wörld")
    assert g["logs"][1] == {"time": 5, "contents": {"Count": "3"}}
    # golden bytes of the smallest group: field 1 {Time=1} , Topic "", Source ""
    assert sls.encode_log_group([(1, [])]) == bytes([0x0A, 0x02, 0x08, 0x01, 0x1A, 0x00, 0x22, 0x00])


def test_sls_event_backend_over_http(monkeypatch):
    """SaveEvent -> signed PutLogs with the reference's fields and retry envelope; ListEvent -> histogram count,
    GetLogs in pages of 100, sorted by first timestamp, one entry per involved object
    (reference pkg/storage/backends/events/aliyun_sls/sls_logstore.go:81-140, 231-257)."""
    from kubedl_b200.plugins.storage import sls
    fake = _FakeSls()
    env = {"SLS_ENDPOINT": fake.url, "SLS_KEY_ID": "kid", "SLS_KEY_SECRET": "s3cret", "SLS_PROJECT": "proj",
           "SLS_LOG_STORE": "store"}
    holds = []
    b = sls.SlsEventBackend(env=env, sleep=holds.append)
    try:
        with pytest.raises(ValueError, match="empty sls key secret"):
            sls.SlsEventBackend(env={**env, "SLS_KEY_SECRET": ""}).initialize()
        b.initialize()
        assert b.name() == "aliyun-sls"
        # retry envelope: a quota error (800 ms hold), a server error (200 ms hold), then success
        fake.fail = ["WriteQuotaExceed", "ServerBusy"]
        b.save_event(_k8s_event(0, "uid-0", "job1-worker-0"), "cn-test")
        assert holds == [0.8, 0.2] and len(fake.logs) == 1
        first = fake.logs[0]
        assert [k for k in first if not k.startswith("__")] == ["FirstTimestamp", "LastTimestamp", "Count", "Name", "Kind",
                                                                "ObjUID", "ObjNamespace", "ObjName", "Type", "Reason",
                                                                "Message", "Region"]
        assert first["Region"] == "cn-test" and first["__source__"] == "pytorch-controller/node-0"
        assert first["__time__"] == "1767225600"
        # a permanent failure is raised after 10 tries
        fake.fail = ["Unauthorized"] * 10
        with pytest.raises(sls.SlsError, match="Unauthorized"):
            b.save_event(_k8s_event(1, "uid-x", "job1-worker-9"), "")
        assert sum(1 for r in fake.requests if r[0] == "POST") == 3 + 10
        # 250 events of job1 (50 pods, 5 events each, written newest first), some of another job and namespace
        for i in reversed(range(1, 251)):
            b.save_event(_k8s_event(i, f"uid-{i % 50}", f"job1-worker-{i % 50}"), "")
        b.save_event(_k8s_event(300, "uid-other", "job2-worker-0"), "")
        b.save_event(_k8s_event(301, "uid-ns", "job1-worker-0", ns="ns2"), "")
        fake.requests.clear()
        got = b.list_event("ns1", "job1", "2026-01-01T00:00:00Z", "2026-01-01T01:00:00Z")
        assert [r[1] for r in fake.requests] == ["histogram", "log", "log", "log"]          # 251 logs -> 3 pages
        assert len(got) == 50 and {e.obj_uid for e in got} == {f"uid-{i}" for i in range(50)}
        assert [e.first_timestamp for e in got] == sorted(e.first_timestamp for e in got)
        assert got[0].obj_uid == "uid-0" and got[0].region == "cn-test" and got[0].count == 1   # the earliest per object
        assert b.list_event("ns1", "job1", "2027-01-01T00:00:00Z", "2027-01-02T00:00:00Z") == []
    finally:
        fake.stop()


def test_registry_picks_the_sls_backend_when_an_endpoint_is_configured(monkeypatch):
    from kubedl_b200.plugins.storage import backends as B
    from kubedl_b200.plugins.storage import sls
    B.REGISTRY.remove_event_backend("aliyun-sls")
    monkeypatch.setenv("SLS_ENDPOINT", "cn-hangzhou.log.aliyuncs.com")
    try:
        B.register_storage_backends()
        assert isinstance(B.REGISTRY.get_event_backend("aliyun-sls"), sls.SlsEventBackend)
    finally:
        B.REGISTRY.remove_event_backend("aliyun-sls")
        monkeypatch.delenv("SLS_ENDPOINT")
        B.register_storage_backends()
    assert isinstance(B.REGISTRY.get_event_backend("aliyun-sls"), B.JsonlEventBackend)
