"""The master-side dispatcher against in-process workers (werkzeug test client as the transport)."""
import json

import pytest
from werkzeug.test import Client

from dli_b200.dispatcher import DispatchError, Node, dispatch_inference
from dli_b200.worker import Worker


class _Resp:
    def __init__(self, r):
        self.status_code, self.text, self._r = r.status_code, r.get_data(as_text=True), r

    def json(self):
        return self._r.get_json()


def _transport(workers, log):
    def post(url, json=None, headers=None, timeout=None):
        base, path = url.rsplit("/", 1)
        log.append((base, "/" + path, json, headers, timeout))
        return _Resp(Client(workers[base]).post("/" + path, json=json, headers=headers or {}))
    return post


def test_dispatch_orders_shards_announces_them_and_sends_auth(tmp_path):
    w = Worker(env={"MODEL_CACHE_DIR": "/nonexistent", "AUTH_ENABLED": "1", "AUTH_KEY": "s3", "USE_GPU": "0"})
    log = []
    post = _transport({"http://a:5000": w}, log)
    p0, p1 = tmp_path / "shard_0", tmp_path / "shard_1"
    p0.mkdir(), p1.mkdir()
    node = Node("http://a:5000", shards={"m": {1: str(p1), 0: str(p0)}})
    # no CUDA here, so the chain itself fails inside the worker — with the worker's own message
    with pytest.raises(DispatchError, match="Inference failed"):
        dispatch_inference([Node("http://dead:5000", is_active=False), node], "m", "1 2 3", auth_key="s3", post=post)
    paths = [(p, j.get("shard_id")) for _, p, j, _, _ in log]
    assert paths == [("/load_shard", 0), ("/load_shard", 1), ("/inference", None)]
    assert log[-1][2]["shard_ids"] == [0, 1] and log[-1][2]["max_length"] == 100 and log[-1][2]["timeout"] == 60
    assert all(h == {"Authorization": "Bearer s3"} for _, _, _, h, _ in log) and log[-1][4] == 120
    assert sorted(w.loaded_shards["m"]) == [0, 1]
    # wrong key -> the worker's 401 surfaces as a dispatch error
    with pytest.raises(DispatchError, match="401"):
        dispatch_inference([node], "m", "1 2 3", auth_key="nope", post=post)


def test_dispatch_without_shards_loads_model_on_first_active_node():
    w = Worker(env={"MODEL_CACHE_DIR": "/nonexistent", "USE_GPU": "0"})
    log = []
    post = _transport({"http://b:5000": w}, log)
    with pytest.raises(DispatchError, match="Failed to load model"):
        dispatch_inference([Node("http://b:5000")], "org/none", "hi", post=post)
    assert [p for _, p, _, _, _ in log] == ["/load_model"] and log[0][4] == 300
    with pytest.raises(DispatchError, match="No active worker nodes available"):
        dispatch_inference([Node("http://b:5000", is_active=False)], "org/none", "hi", post=post)


def test_dispatch_with_one_shard_per_node_addresses_the_pipeline_head(tmp_path):
    """The reference's deployment: shard i on node i (models.py:27-28).  Every node gets its own /load_shard, and the
    node of the FIRST shard gets one /inference naming all shards in order (pipeline_worker.py serves it)."""
    ws = {f"http://n{i}:5000": Worker(env={"MODEL_CACHE_DIR": "/nonexistent", "USE_GPU": "0"}) for i in range(3)}
    log = []
    post = _transport(ws, log)
    nodes = []
    for i in (2, 0, 1):  # registered in no particular order
        p = tmp_path / f"shard_{i}"
        p.mkdir()
        nodes.append(Node(f"http://n{i}:5000", shards={"m": {i: str(p)}}))
    with pytest.raises(DispatchError, match="Inference failed"):  # no CUDA here: the worker's own message
        dispatch_inference(nodes, "m", "1 2 3", post=post)
    assert [(b, p, j.get("shard_id")) for b, p, j, _, _ in log[:3]] == [
        ("http://n0:5000", "/load_shard", 0), ("http://n1:5000", "/load_shard", 1), ("http://n2:5000", "/load_shard", 2)]
    assert log[3][0] == "http://n0:5000" and log[3][1] == "/inference" and log[3][2]["shard_ids"] == [0, 1, 2]
    assert len(log) == 4
    assert all(list(ws[f"http://n{i}:5000"].loaded_shards["m"]) == [i] for i in range(3))
