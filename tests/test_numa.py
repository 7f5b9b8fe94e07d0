"""utils/numa.py on a fabricated sysfs tree: node of the GPU, CPU list parsing, best-effort
behaviour (one node, missing files, PSB_NUMA_BIND=0, empty intersection)."""
import os
import types

import pytest
import torch

from parameter_server_distributed_b200.utils import numa


def _sysfs(tmp_path, nodes, gpu_node):
    for n, cpus in nodes.items():
        d = tmp_path / "devices/system/node" / f"node{n}"
        d.mkdir(parents=True)
        (d / "cpulist").write_text(cpus + "\n")
    (tmp_path / "devices/system/node/online").write_text("0\n")
    dev = tmp_path / "bus/pci/devices/0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text(f"{gpu_node}\n")
    return str(tmp_path)


@pytest.fixture()
def fake_gpu(monkeypatch):
    props = types.SimpleNamespace(pci_domain_id=0, pci_bus_id=0x1B, pci_device_id=0)
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda i: props)
    calls = []
    monkeypatch.setattr(os, "sched_setaffinity", lambda pid, cpus: calls.append(set(cpus)))
    monkeypatch.setattr(numa, "_prefer_node_memory", lambda node: True)
    return calls


def test_cpulist_parsing():
    assert numa._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa._parse_cpulist("") == set()


def test_binds_to_the_cpus_of_the_gpus_node(tmp_path, fake_gpu, monkeypatch):
    allowed = sorted(os.sched_getaffinity(0))
    half = max(1, len(allowed) // 2)
    a, b = allowed[:half], allowed[half:] or allowed[:1]
    fmt = lambda c: ",".join(map(str, c))  # noqa: E731
    root = _sysfs(tmp_path, {0: fmt(a), 1: fmt(b)}, gpu_node=1)
    assert numa.gpu_numa_node(0, root) == 1
    info = numa.bind_to_gpu_node(0, root)
    assert info == {"node": 1, "cpus": len(set(b)), "mempolicy": True}
    assert fake_gpu == [set(b)]


def test_best_effort_cases_leave_the_process_alone(tmp_path, fake_gpu, monkeypatch):
    one = _sysfs(tmp_path / "one", {0: "0-63"}, gpu_node=0)
    assert numa.bind_to_gpu_node(0, one) is None                      # a single node
    unknown = _sysfs(tmp_path / "unknown", {0: "0-3", 1: "4-7"}, gpu_node=-1)
    assert numa.bind_to_gpu_node(0, unknown) is None                  # the platform does not say
    nowhere = _sysfs(tmp_path / "nowhere", {0: "100000-100003", 1: "100004-100007"}, gpu_node=1)
    assert numa.bind_to_gpu_node(0, nowhere) is None                  # cpuset does not intersect
    assert numa.bind_to_gpu_node(0, str(tmp_path / "missing")) is None
    if len(os.sched_getaffinity(0)) >= 4:
        first = str(min(os.sched_getaffinity(0)))
        sliver = _sysfs(tmp_path / "sliver", {0: "100000-100003", 1: first}, gpu_node=1)
        assert numa.bind_to_gpu_node(0, sliver) is None              # one CPU of that node allowed
    two = _sysfs(tmp_path / "two", {0: "0-3", 1: "0-7"}, gpu_node=1)
    monkeypatch.setenv("PSB_NUMA_BIND", "0")
    assert numa.bind_to_gpu_node(0, two) is None
    assert fake_gpu == []


def test_real_box_never_raises():
    # no GPU here: properties fail, so nothing is done — and nothing is thrown
    assert numa.gpu_numa_node(0) is None or isinstance(numa.gpu_numa_node(0), int)
    numa.bind_to_gpu_node(0)
    numa._prefer_node_memory(0)
