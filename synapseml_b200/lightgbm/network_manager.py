"""Rendezvous: who are the LightGBM "machines" and on which ports.

Mirror of lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/NetworkManager.scala: the driver collects one
"host:port" line per main worker and answers with the machine list (:322-328, :384-445); each main worker then calls
LGBM_NetworkInit(machines, localPort, 120, n) with 3 retries and exponential backoff (:195-218).  Under torchrun the
ranks are the executors; the collection step uses torch.distributed's object all-gather (gloo on CPU) instead of the
driver ServerSocket.  With one machine nothing is initialised (the native learner is serial).
"""
import os
import socket
import time

from .native import lightgbmlib, validate

DEFAULT_LOCAL_LISTEN_PORT = 12400
DEFAULT_LISTEN_TIMEOUT = 120
NETWORK_RETRIES = 3
INITIAL_DELAY_MS = 1000


def find_open_port(default_listen_port, worker_id, tasks_per_exec=1, max_tries=1000):
    """NetworkManager.findOpenPort (:241-271): scan upwards from basePort + workerId * tasksPerExec."""
    base = default_listen_port + worker_id * tasks_per_exec
    for port in range(base, base + max_tries):
        s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        try:
            s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            s.bind(("", port))
            return port
        except OSError:
            continue
        finally:
            s.close()
    raise RuntimeError("Could not find an open port in [%d, %d)" % (base, base + max_tries))


class NetworkTopologyInfo:
    def __init__(self, machines, local_listen_port, rank, world):
        self.lightgbm_network_string = machines
        self.local_listen_port = local_listen_port
        self.rank, self.world = rank, world


def get_global_network_info(default_listen_port=DEFAULT_LOCAL_LISTEN_PORT, group=None):
    """One entry per rank, ordered by rank (the reference orders machines by their minimum partition id, :322-328)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return NetworkTopologyInfo("", default_listen_port, 0, 1)
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    port = find_open_port(default_listen_port, rank)
    host = os.environ.get("LIGHTGBM_B200_HOST", "127.0.0.1")
    entries = [None] * world
    dist.all_gather_object(entries, "%s:%d" % (host, port), group=group)
    return NetworkTopologyInfo(",".join(entries), port, rank, world)


def init_lightgbm_network(info, timeout=DEFAULT_LISTEN_TIMEOUT, retries=NETWORK_RETRIES, delay_ms=INITIAL_DELAY_MS):
    """NetworkManager.initLightGBMNetwork (:195-218)."""
    if info.world <= 1:
        return
    for attempt in range(retries + 1):
        rc = lightgbmlib().LGBM_NetworkInit(info.lightgbm_network_string.encode(), info.local_listen_port, timeout, info.world)
        if rc == 0:
            return
        if attempt == retries:
            validate(rc, "Network init")
        time.sleep(delay_ms / 1000.0)
        delay_ms *= 2


def free_lightgbm_network():
    """BasePartitionTask.cleanup (BasePartitionTask.scala:278-288)."""
    validate(lightgbmlib().LGBM_NetworkFree(), "Finalize network")
