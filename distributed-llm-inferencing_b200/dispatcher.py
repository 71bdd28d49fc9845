"""Client-side dispatch of an inference request to worker nodes — the caller of the worker
boundary (reference master/dashboard/views.py:305-455 `process_inference_request`), without
Django: plain functions over a `post(url, json, headers, timeout)` callable (requests.post by
default).  SURVEY.md §8f rank 2.  Kept semantics: shards known -> /inference with `shard_ids`
(max_length 100, timeout 60, HTTP timeout 120, views.py:345-355); no shards -> /load_model
(HTTP timeout 300) then /inference on the first active node (views.py:391-421); a non-200
answer or `status != success` is a failure with the worker's message.
Fixed relative to the reference: ALL shard ids of the model are sent, ordered, to the pipeline
head — the node holding the first shard — whether the shards sit on one node or one per node (the
reference takes "the first node that has shards" and sends only its ids, views.py:337-340); shards that
are registered with the master but not yet on the worker are (re)announced with /load_shard
first (no reference view ever calls it); the Authorization header is sent when a key is
configured (the reference master never sends it, so AUTH_ENABLED=1 breaks it).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional


@dataclass
class Node:
    url: str  # e.g. http://10.0.0.5:5000
    is_active: bool = True
    # shard_id -> shard_path of the shards this node holds for a model: {model: {id: path}}
    shards: Dict[str, Dict[object, str]] = field(default_factory=dict)


class DispatchError(RuntimeError):
    pass


def _default_post(url, json=None, headers=None, timeout=None):
    import requests

    return requests.post(url, json=json, headers=headers, timeout=timeout)


def _check(resp, what):
    if resp.status_code != 200:
        raise DispatchError(f"{what}: node returned status code {resp.status_code}: {resp.text}")
    data = resp.json()
    if data.get("status") != "success":
        raise DispatchError(f"{what}: {data.get('message', 'Unknown error')}")
    return data


def dispatch_inference(nodes: List[Node], model_name: str, prompt: str, max_length: int = 100, timeout: int = 60,
                       auth_key: Optional[str] = None, post: Callable = _default_post, extra: Optional[dict] = None) -> str:
    """Returns the generated text, or raises DispatchError with the failure message the
    reference would store in InferenceRequest.error."""
    headers = {"Authorization": f"Bearer {auth_key}"} if auth_key else {}
    active = [n for n in nodes if n.is_active]
    holders = [n for n in active if n.shards.get(model_name)]
    body = {"model_name": model_name, "prompt": prompt, "max_length": max_length, "timeout": timeout, **(extra or {})}
    if holders:
        order = lambda i: (str(type(i)), i)  # noqa: E731 — ids of one model are all ints or all strings
        # every shard is announced to the node that holds it (one node per (model, shard):
        # master/dashboard/models.py:27-28), lowest id first
        owner = {}
        for n in holders:
            for sid in n.shards[model_name]:
                owner.setdefault(sid, n)
        ids = sorted(owner, key=order)
        for sid in ids:
            n = owner[sid]
            _check(post(f"{n.url}/load_shard", json={"model_name": model_name, "shard_id": sid,
                                                     "shard_path": n.shards[model_name][sid]},
                        headers=headers, timeout=300), f"load_shard {sid}")
        # ONE /inference with ALL shard ids, in order, to the pipeline head: the node holding the first shard.  With
        # every shard on one node that node chains them in-process; with one shard per node (the reference's
        # deployment, served here by pipeline_worker.py) the head drives the group.  The reference instead takes
        # "the first node that has shards" and sends only that node's ids (views.py:337-340).
        head = owner[ids[0]]
        return _check(post(f"{head.url}/inference", json={**body, "shard_ids": ids}, headers=headers, timeout=120),
                      "inference")["result"]
    if not active:
        raise DispatchError("No active worker nodes available")
    node = active[0]
    _check(post(f"{node.url}/load_model", json={"model_name": model_name}, headers=headers, timeout=300),
           "Failed to load model")
    return _check(post(f"{node.url}/inference", json=body, headers=headers, timeout=120), "inference")["result"]
