"""Host-side NUMA placement helper (shellac_b200/hostmem.py): pure sysfs parsing, no GPU needed."""
import os

from shellac_b200 import hostmem


def test_cpulist_forms():
    assert hostmem.parse_cpulist('0-3,8,10-11') == [0, 1, 2, 3, 8, 10, 11]
    assert hostmem.parse_cpulist('5') == [5]
    assert hostmem.parse_cpulist('') == [] and hostmem.parse_cpulist(None) == []


def test_node_lookup_on_a_fake_sysfs(tmp_path, monkeypatch):
    dev = tmp_path / 'bus/pci/devices/0000:1b:00.0'
    dev.mkdir(parents=True)
    (dev / 'numa_node').write_text('1\n')
    node = tmp_path / 'devices/system/node/node1'
    node.mkdir(parents=True)
    mine = sorted(os.sched_getaffinity(0))
    (node / 'cpulist').write_text('%d\n' % mine[0])
    monkeypatch.setattr(hostmem, 'gpu_pci_address', lambda d: '0000:1b:00.0')
    assert hostmem.gpu_numa_node(0, str(tmp_path)) == 1
    assert hostmem.node_cpus(1, str(tmp_path)) == [mine[0]]
    before = os.sched_getaffinity(0)
    try:
        assert hostmem.bind_near_gpu(0, str(tmp_path)) == (1, 1)
        assert os.sched_getaffinity(0) == {mine[0]}
    finally:
        os.sched_setaffinity(0, before)


def test_vm_without_numa_info_is_left_alone(tmp_path, monkeypatch):
    dev = tmp_path / 'bus/pci/devices/0000:1b:00.0'
    dev.mkdir(parents=True)
    (dev / 'numa_node').write_text('-1\n')
    monkeypatch.setattr(hostmem, 'gpu_pci_address', lambda d: '0000:1b:00.0')
    before = os.sched_getaffinity(0)
    assert hostmem.bind_near_gpu(0, str(tmp_path)) is None
    assert os.sched_getaffinity(0) == before
