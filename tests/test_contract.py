"""Operator-facing contract: TFJob replica specs in, TF_CONFIG per task out (bit-exact strings).

The reference only names the files that own this contract (images/tf3.PNG: v1alpha1/types.go,
defaults.go, validation/validation.go -- each with a *_test.go beside it -- and
trainer/replicas.go); these tests play the role of those defaults/validation unit tests."""
import glob
import json
import os

import pytest

import b200ps
from b200ps import B200PSError, Cluster

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "tf_config_*.json"))))
def test_tf_config_golden_strings(path):
    fx = json.load(open(path))
    c = Cluster.from_tfjob(fx["tfjob"])
    for key, want in fx["tf_config"].items():
        job, idx = key.split(":")
        got = c.tf_config_for(job, int(idx))
        assert got == want
        # and it is valid JSON with the fields a TF task reads
        d = json.loads(got)
        assert d["task"] == {"type": job, "index": int(idx)} and d["environment"] == "cloud"
        # round trip through the parser
        c2, t, i = Cluster.from_tf_config(got)
        assert (t, i) == (job, int(idx)) and c2.as_dict() == c.as_dict()
        assert c2.tf_config_for(job, int(idx)) == got


def test_defaults_replicas_port_type():
    # defaults.go: replicas -> 1, tfPort -> 2222, replica type -> MASTER
    c = Cluster.from_tfjob({"metadata": {"name": "j"}, "spec": {"runtimeId": "r1", "replicaSpecs": [{}, {"tfReplicaType": "PS"}]}})
    assert c.as_dict() == {"master": ["j-master-r1-0:2222"], "ps": ["j-ps-r1-0:2222"]}
    assert c.num_workers == 1 and c.num_shards == 1 and c.worker_rank("master", 0) == 0


def test_yaml_input_and_map_form():
    y = """
apiVersion: kubeflow.org/v1alpha2
kind: TFJob
metadata:
  name: wide-deep
spec:
  tfReplicaSpecs:
    Chief: {replicas: 1}
    PS: {replicas: 8}
    Worker: {replicas: 7}
"""
    c = Cluster.from_tfjob(y)
    assert c.num_shards == 8 and c.num_workers == 8
    assert c.task_address("worker", 6) == "wide-deep-worker-6:2222"
    assert c.worker_rank("chief", 0) == 0 and c.worker_rank("worker", 0) == 1
    cfg = json.loads(c.tf_config_for("chief", 0))
    assert list(cfg["cluster"]) == ["chief", "ps", "worker"]  # sorted keys, as Go marshals a map


def test_runtime_id_is_deterministic_without_one():
    spec = {"metadata": {"name": "abc"}, "spec": {"replicaSpecs": [{"tfReplicaType": "WORKER", "replicas": 2}]}}
    a, b = Cluster.from_tfjob(spec), Cluster.from_tfjob(json.dumps(spec))
    assert a.as_dict() == b.as_dict()
    host = a.task_address("worker", 1)
    assert host.startswith("abc-worker-") and host.endswith("-1:2222") and len(host.split("-")[2]) == 4


@pytest.mark.parametrize("spec,needle", [
    ({"spec": {"replicaSpecs": [{"tfReplicaType": "EVALUATOR"}]}}, "must be one of MASTER, PS, WORKER"),
    ({"spec": {"replicaSpecs": [{"tfReplicaType": "MASTER", "replicas": 2}]}}, "must have replicas = 1"),
    ({"spec": {"replicaSpecs": [{"tfReplicaType": "WORKER", "replicas": 0}]}}, "replicas must be an integer >= 1"),
    ({"spec": {"replicaSpecs": [{"tfReplicaType": "WORKER", "tfPort": 70000}]}}, "tfPort must be in 1..65535"),
    ({"spec": {"replicaSpecs": [{"tfReplicaType": "PS"}, {"tfReplicaType": "PS"}]}}, "appears twice"),
    ({"spec": {"replicaSpecs": [{"tfReplicaType": "PS"}]}}, "at least one MASTER/Chief or WORKER"),
    ({"spec": {"replicaSpecs": []}}, "no replica specs"),
    ({"spec": {}}, "neither replicaSpecs nor tfReplicaSpecs"),
    ({}, "no spec"),
    ({"metadata": {"name": "Bad_Name"}, "spec": {"replicaSpecs": [{}]}}, "DNS-1123"),
    ({"spec": {"tfReplicaSpecs": {"Chief": {}, "Master": {}}}}, "not both"),
])
def test_validation_rejects(spec, needle):
    with pytest.raises(B200PSError) as ei:
        Cluster.from_tfjob(spec)
    assert ei.value.status == b200ps._lib.ERR_VALIDATION and needle in str(ei.value)


@pytest.mark.parametrize("text", ["", "{", "[1,2", '{"spec": tru}', '{"a":"\\q"}', '{"a":1} x'])
def test_malformed_json_is_a_parse_error(text):
    with pytest.raises(B200PSError) as ei:
        Cluster.from_tfjob(text if text.startswith(("{", "[")) else "{" + text)
    assert ei.value.status in (b200ps._lib.ERR_PARSE, b200ps._lib.ERR_VALIDATION)


def test_tf_config_errors():
    good = '{"cluster":{"ps":["a:1"],"worker":["b:1","c:1"]},"task":{"type":"worker","index":1}}'
    c, t, i = Cluster.from_tf_config(good)
    assert (t, i, c.num_shards, c.num_workers) == ("worker", 1, 1, 2)
    with pytest.raises(B200PSError, match="not in the cluster spec"):
        Cluster.from_tf_config(good.replace('"index":1', '"index":2'))
    with pytest.raises(B200PSError, match="no cluster object"):
        Cluster.from_tf_config('{"task":{"type":"ps","index":0}}')
    with pytest.raises(B200PSError, match="no task"):
        c.tf_config_for("ps", 5)
    with pytest.raises(B200PSError):
        c.worker_rank("ps", 0)


def test_json_escaping_matches_go():
    # Go's json.Marshal escapes <, >, & and control characters; hosts from a hand-written TF_CONFIG may hold them
    c, _, _ = Cluster.from_tf_config('{"cluster":{"worker":["a<b>&\\"c\\\\:1"]},"task":{"type":"worker","index":0}}')
    s = c.tf_config_for("worker", 0)
    assert '"a\\u003cb\\u003e\\u0026\\"c\\\\:1"' in s
    assert json.loads(s)["cluster"]["worker"] == ['a<b>&"c\\:1']


def test_placement():
    # shard s on GPU s mod G, worker w on GPU w mod G
    assert [b200ps.placement_gpu(w, 4) for w in range(6)] == [0, 1, 2, 3, 0, 1]
    assert [b200ps.placement_gpu(s, 1) for s in range(3)] == [0, 0, 0]
    with pytest.raises(B200PSError):
        b200ps.placement_gpu(0, 0)


def test_json_numbers_and_escapes_follow_the_grammar():
    """Advisor finding (round 1): numbers go through the JSON grammar before strtod (no 0x10 / inf / nan /
    leading '+' / '01'), out-of-range doubles are rejected before any integer cast, and a lone or
    mismatched UTF-16 surrogate in a \\u escape is a parse error instead of invalid UTF-8 in TF_CONFIG."""
    def job(rep):
        return '{"metadata":{"name":"a"},"spec":{"replicaSpecs":[{"replicas":%s,"tfReplicaType":"WORKER"}]}}' % rep

    for bad, status in (("0x10", b200ps._lib.ERR_PARSE), ("-inf", b200ps._lib.ERR_PARSE), ("+1", b200ps._lib.ERR_PARSE),
                        ("01", b200ps._lib.ERR_PARSE), ("1.", b200ps._lib.ERR_PARSE), (".5", b200ps._lib.ERR_PARSE),
                        ("1e", b200ps._lib.ERR_PARSE), ("1e999", b200ps._lib.ERR_VALIDATION), ("1e300", b200ps._lib.ERR_VALIDATION),
                        ("2.5", b200ps._lib.ERR_VALIDATION)):
        with pytest.raises(b200ps.B200PSError) as ei:
            b200ps.Cluster.from_tfjob(job(bad))
        assert ei.value.status == status, bad
    assert [b200ps.Cluster.from_tfjob(job(ok)).num_workers for ok in ("2", "2.0", "2e0", "20e-1")] == [2, 2, 2, 2]
    for name in ('"\\ud800x"', '"\\ud800\\u0041"', '"\\udc00"'):
        with pytest.raises(b200ps.B200PSError) as ei:
            b200ps.Cluster.from_tfjob('{"metadata":{"name":%s},"spec":{"replicaSpecs":[{"replicas":1,"tfReplicaType":"WORKER"}]}}' % name)
        assert ei.value.status == b200ps._lib.ERR_PARSE
