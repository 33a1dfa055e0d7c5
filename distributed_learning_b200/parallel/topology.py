"""GPU / NVLink topology discovery.

The reference picks its transport by naming a network interface (``i`` -> ``GLOO_SOCKET_IFNAME``,
``src/main.py:306-308``).  On an NVSwitch box the question is different: which GPUs can map
each other's memory, do they support native atomics over the link, how many NVLink links are up,
and can the switch reduce (multicast / NVLS).  ``query_topology`` asks the CUDA runtime, the
driver and NVML (``csrc/runtime.cpp: dlb_topology_json``) and returns a dict;
``describe_topology`` renders it for the log; ``all_to_all_capable`` is the gate the native
backend checks before it maps peers.
"""
from __future__ import annotations

import ctypes
import json
from typing import Dict, Sequence


def query_topology() -> Dict:
    from ..ops._lib import check, lib

    buf = ctypes.create_string_buffer(1 << 16)
    check(lib().dlb_topology_json(buf, len(buf)))
    return json.loads(buf.value.decode())


def all_to_all_capable(topo: Dict, devices: Sequence[int]) -> bool:
    """True when every listed device can map every other one's memory."""
    pa = topo.get("peer_access", [])
    return all(pa[a][b] for a in devices for b in devices if a != b) if pa else False


def describe_topology(topo: Dict = None) -> str:
    topo = topo if topo is not None else query_topology()
    devs = topo.get("devices", [])
    lines = [f"topology: {len(devs)} GPU(s), CUDA runtime {topo.get('runtime')}, driver {topo.get('driver')}, "
             f"NVML {'yes' if topo.get('nvml') else 'no'}"]
    for d in devs:
        links = d["nvlinks_active"]
        lines.append(f"  gpu{d['index']}: {d['name']} sm_{d['cc'].replace('.', '')} {d['sms']} SMs "
                     f"{d['mem_gib']:.0f} GiB L2 {d['l2_mib']:.0f} MiB pci {d['pci']} "
                     f"nvlink {'?' if links < 0 else links} links v{d['nvlink_version']} "
                     f"multicast {'yes' if d['multicast'] else 'no'}")
    pa = topo.get("peer_access", [])
    if len(pa) > 1:
        full = all(all(r) for r in pa)
        lines.append(f"  peer access: {'all-to-all' if full else 'partial'}")
        if not full:
            lines += ["    " + " ".join(str(x) for x in row) for row in pa]
        na = topo.get("native_atomics", [])
        if na:
            lines.append(f"  native atomics over links: {'all pairs' if all(all(r) for r in na) else 'partial'}")
    return "\n".join(lines)
