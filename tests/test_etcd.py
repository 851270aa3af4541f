"""etcd-compatible store: core semantics, the JSON gateway + client over real
sockets, and the controller's np sync against it (tier T-etcd)."""
import threading
import time

import pytest

from paddle_operator_b200.api.types import PaddleJob
from paddle_operator_b200.controller.elastic import np_key, sync_np
from paddle_operator_b200.runtime.etcd_client import EtcdClient
from paddle_operator_b200.runtime.etcd_server import EtcdServer
from paddle_operator_b200.runtime.etcd_store import EtcdStore, InProcessEtcd, prefix_range_end


def test_store_revisions_ranges_and_txn():
    s = EtcdStore()
    try:
        assert prefix_range_end(b"/paddle/") == b"/paddle0" and prefix_range_end(b"\xff") == b"\0"
        _, r1 = s.put("/a/1", "x")
        _, r2 = s.put("/a/2", "y")
        assert r2 == r1 + 1
        prev, _ = s.put("/a/1", "z", prev_kv=True)
        assert prev.value == b"x"
        kvs, count, _ = s.range("/a/", prefix_range_end(b"/a/"))
        assert [k.key for k in kvs] == [b"/a/1", b"/a/2"] and count == 2
        assert kvs[0].version == 2 and kvs[0].create_revision == r1 and kvs[0].mod_revision > r2
        ok, _, _ = s.txn([("/a/3", "create", "=", 0)], [("put", "/a/3", "new")], [])
        assert ok
        ok, res, _ = s.txn([("/a/3", "create", "=", 0)], [("put", "/a/3", "again")], [("range", "/a/3")])
        assert not ok and res[0][1][0].value == b"new"
        ok, _, _ = s.txn([("/a/3", "value", "=", "new")], [("delete", "/a/", prefix_range_end(b"/a/"))], [])
        assert ok and s.range("/a/", prefix_range_end(b"/a/"))[1] == 0
    finally:
        s.close()


def test_lease_expiry_and_watch():
    s = EtcdStore(reap_interval=0.02)
    try:
        events = []
        s.watch("/n/", prefix_range_end(b"/n/"), fn=lambda evs: events.extend((e.type, e.kv.key) for e in evs))
        lid, ttl = s.lease_grant(1)
        s.put("/n/a", "1", lease=lid)
        s.put("/n/b", "1", lease=lid)
        s.put("/other", "1")
        for _ in range(4):  # keep-alive holds it past its TTL
            time.sleep(0.4)
            assert s.lease_keepalive(lid) == ttl
        assert s.range("/n/", prefix_range_end(b"/n/"))[1] == 2
        time.sleep(1.4)  # no keep-alive: the lease expires and takes its keys along
        assert s.range("/n/", prefix_range_end(b"/n/"))[1] == 0 and s.lease_keepalive(lid) == 0
        assert events == [("PUT", b"/n/a"), ("PUT", b"/n/b"), ("DELETE", b"/n/a"), ("DELETE", b"/n/b")]
        # replay from a revision
        seen = []
        s.watch("/n/", prefix_range_end(b"/n/"), start_revision=1, fn=lambda evs: seen.extend(e.type for e in evs))
        assert seen == ["PUT", "PUT", "DELETE", "DELETE"]
    finally:
        s.close()


@pytest.fixture
def server():
    with EtcdServer("127.0.0.1", 0) as srv:
        yield srv


def test_gateway_client_roundtrip(server):
    c = EtcdClient(server.endpoint)
    assert c.ping() and c.endpoints() == [server.endpoint]
    c.put("/paddle/default-job/np", "3")
    assert c.get("/paddle/default-job/np") == [("/paddle/default-job/np", "3")]
    assert c.get("/missing") == []
    assert c.put_if_absent("/paddle/default-job/np", "9") is False
    assert c.put_if_absent("/paddle/default-job/x", "1") is True
    lid = c.lease_grant(5)
    c.put("/paddle/default-job/nodes/10.0.0.1", "a", lease=lid)
    c.put("/paddle/default-job/nodes/10.0.0.2", "b", lease=lid)
    assert [k for k, _ in c.get_prefix("/paddle/default-job/nodes/")] == ["/paddle/default-job/nodes/10.0.0.1",
                                                                           "/paddle/default-job/nodes/10.0.0.2"]
    assert c.lease_keepalive(lid) == 5
    got, ev = [], threading.Event()

    def on(evs):
        got.extend(evs)
        ev.set()

    cancel = c.watch_prefix("/paddle/default-job/", on)
    c.put("/paddle/default-job/np", "4")
    assert ev.wait(5)
    assert ("PUT", "/paddle/default-job/np", "4") in got
    ev.clear()
    c.lease_revoke(lid)
    deadline = time.time() + 5
    while time.time() < deadline and sum(1 for t, *_ in got if t == "DELETE") < 2:
        time.sleep(0.02)
    assert sum(1 for t, *_ in got if t == "DELETE") == 2
    cancel()
    assert c.delete_prefix("/paddle/") == 2 and c.get_prefix("/paddle/") == []
    assert c.lease_keepalive(12345) == 0


@pytest.mark.parametrize("kind", ["inproc", "http"])
def test_controller_np_sync_against_the_store(server, kind):
    etcd = InProcessEtcd() if kind == "inproc" else EtcdClient(server.endpoint)
    job = PaddleJob.from_dict({"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob",
                               "metadata": {"name": "resnet-elastic", "namespace": "default"},
                               "spec": {"elastic": 1, "worker": {"replicas": 4, "template": {}}}})
    job.status.mode = "Collective"
    key = np_key("default", "resnet-elastic")
    assert key == "/paddle/default-resnet-elastic/np"
    assert sync_np(etcd, job) is None and etcd.get(key) == []  # the controller never creates the key
    etcd.put(key, "3")
    assert sync_np(etcd, job) == "4" and etcd.get(key) == [(key, "4")]
    assert sync_np(etcd, job) is None  # already in sync
    from paddle_operator_b200.api.types import ResourceSpec

    job.status.mode = "PS"
    job.spec.ps = ResourceSpec(replicas=2)
    job.spec.worker.replicas = 8
    assert sync_np(etcd, job) is None and etcd.get(key) == [(key, "4")]  # never for PS jobs
    # an elastic job scaled to ONE worker reports mode Single; np must still follow (see controller/elastic.py)
    job.spec.ps = None
    job.status.mode = "Single"
    job.spec.worker.replicas = 1
    assert sync_np(etcd, job) == "1"
