"""Scratch probe (torchrun, one rank per GPU): does binding each rank to the CPUs of its GPU's NUMA node change the
device -> pinned-host bandwidth when all ranks read back at once?  Prints one line per rank per mode."""
import os, sys, time, glob
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank = int(os.environ.get('RANK', 0)); world = int(os.environ.get('WORLD_SIZE', 1)); lr = int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', lr))


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def allmax(v):
    t = torch.tensor([v], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gpu_numa(idx):
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(idx)
    bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
    if isinstance(bdf, bytes):
        bdf = bdf.decode()
    bdf = bdf.lower()
    cands = [bdf, bdf[4:] if len(bdf) > 12 else bdf]
    node = None
    for c in cands:
        p = f'/sys/bus/pci/devices/{c}/numa_node'
        if os.path.exists(p):
            node = int(open(p).read().strip())
            break
    try:
        n_words = (os.cpu_count() + 63) // 64
        aff = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = [64 * w + b for w, m in enumerate(aff) for b in range(64) if (m >> b) & 1]
    except Exception as e:   # noqa
        cpus = None
    return bdf, node, cpus


def parse_cpulist(s):
    out = []
    for part in s.strip().split(','):
        if not part:
            continue
        if '-' in part:
            a, b = part.split('-'); out += list(range(int(a), int(b) + 1))
        else:
            out.append(int(part))
    return out


def measure(tag, nbytes=402653184, reps=4):
    buf = torch.empty(nbytes, dtype=torch.uint8, device='cuda')
    hbuf = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    hbuf.copy_(buf); torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        hbuf.copy_(buf, non_blocking=True)
        torch.cuda.synchronize()
    mine = 1e3 * (time.perf_counter() - t0) / reps
    worst = allmax(mine)
    # host -> device too
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        buf.copy_(hbuf, non_blocking=True)
        torch.cuda.synchronize()
    h2d = allmax(1e3 * (time.perf_counter() - t0) / reps)
    # alone (ranks take turns)
    alone = None
    for r in range(world):
        barrier()
        if r == rank:
            t0 = time.perf_counter()
            for _ in range(2):
                hbuf.copy_(buf, non_blocking=True)
                torch.cuda.synchronize()
            alone = 1e3 * (time.perf_counter() - t0) / 2
    barrier()
    print(f'[{tag}] rank {rank}: d2h all-at-once {mine:7.2f} ms (max over ranks {worst:7.2f}; {nbytes / worst / 1e6:6.1f} GB/s/rank)  '
          f'h2d all-at-once max {h2d:7.2f} ms   d2h alone {alone:7.2f} ms ({nbytes / alone / 1e6:5.1f} GB/s)', flush=True)
    del buf, hbuf


bdf, node, cpus = gpu_numa(lr)
nodes = sorted(glob.glob('/sys/devices/system/node/node[0-9]*'))
print(f'rank {rank}: gpu {lr} bdf {bdf} sysfs numa_node {node} nvml cpu affinity '
      f'{(cpus[0], cpus[-1], len(cpus)) if cpus else None}; process affinity {len(os.sched_getaffinity(0))} cpus; '
      f'{len(nodes)} numa nodes', flush=True)
if rank == 0:
    for n in nodes:
        print('  ', os.path.basename(n), open(n + '/cpulist').read().strip(), flush=True)
    os.system('nvidia-smi topo -m 2>&1 | head -24')
measure('default')
target = None
if node is not None and node >= 0 and os.path.exists(f'/sys/devices/system/node/node{node}/cpulist'):
    target = parse_cpulist(open(f'/sys/devices/system/node/node{node}/cpulist').read())
elif cpus:
    target = cpus
if target:
    allowed = os.sched_getaffinity(0)
    target = [c for c in target if c in allowed]
if target:
    os.sched_setaffinity(0, target)
    print(f'rank {rank}: bound to {len(target)} cpus ({target[0]}..{target[-1]})', flush=True)
    measure('bound')
else:
    print(f'rank {rank}: no NUMA information: not bound', flush=True)
if world > 1:
    dist.destroy_process_group()
