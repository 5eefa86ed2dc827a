"""Where the end-to-end loss step spends its host time (diagnostics).  Runs bench.py's end-to-end step (pinned host
embeddings -> device, SupConLoss(graph=True), loss back to the host, sync) in a few variants and prints wall time per
step plus the host time of each segment.   python tools/prof_host_loss.py [steps]"""
import os
import sys
import time

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from simclr_pytorch_distributed_b200 import SupConLoss   # noqa: E402
from cuda.bindings import runtime as rt                  # noqa: E402

STEPS = int(sys.argv[1]) if len(sys.argv) > 1 else 400
B, V, D, POOL = 4096, 2, 128, 8
dtype = torch.bfloat16
g = torch.Generator().manual_seed(1)
host = [F.normalize(torch.randn(B, V, D, generator=g), dim=-1).to(dtype).pin_memory() for _ in range(POOL)]
nbytes = host[0].numel() * host[0].element_size()
loss_host = torch.empty((), dtype=dtype).pin_memory()
copy_stream = torch.cuda.Stream()
slots = [torch.empty(B, V, D, dtype=dtype, device='cuda') for _ in range(2)]
ready = [torch.cuda.Event() for _ in range(2)]
freed = [torch.cuda.Event() for _ in range(2)]
ones = torch.ones((), dtype=dtype, device='cuda')
crit = SupConLoss(temperature=0.5, engine='bf16', graph=True)
seg = {}
leaves = [s.detach().requires_grad_(True) for s in slots]
loss_read = torch.cuda.Event()


def lap(name, t0):
    t1 = time.perf_counter_ns()
    seg[name] = seg.get(name, 0) + (t1 - t0)
    return t1


def prefetch_torch(j):
    with torch.cuda.stream(copy_stream):
        copy_stream.wait_event(freed[j % 2])
        slots[j % 2].copy_(host[j % POOL], non_blocking=True)
        ready[j % 2].record(copy_stream)


CS = copy_stream.cuda_stream
H2D = rt.cudaMemcpyKind.cudaMemcpyHostToDevice
D2H = rt.cudaMemcpyKind.cudaMemcpyDeviceToHost
host_ptr = [h.data_ptr() for h in host]
slot_ptr = [s.data_ptr() for s in slots]
freed_h = ready_h = None


def prefetch_raw(j):
    rt.cudaStreamWaitEvent(CS, freed_h[j % 2], 0)
    rt.cudaMemcpyAsync(slot_ptr[j % 2], host_ptr[j % POOL], nbytes, H2D, CS)
    rt.cudaEventRecord(ready_h[j % 2], CS)


def variant(name, order_d2h_first, use_ones, raw_copies, no_autograd=False, lean=False):
    global freed_h, ready_h
    cur = torch.cuda.current_stream()
    CUR = cur.cuda_stream
    for e in freed + ready:
        e.record()
    torch.cuda.synchronize()
    freed_h = [e.cuda_event for e in freed]
    ready_h = [e.cuda_event for e in ready]
    prefetch = prefetch_raw if raw_copies else prefetch_torch
    lh_ptr = loss_host.data_ptr()
    entry = None

    def d2h(loss):
        if raw_copies:
            rt.cudaMemcpyAsync(lh_ptr, loss.data_ptr(), 2, D2H, CUR)
        else:
            loss_host.copy_(loss.detach(), non_blocking=True)

    def step(i, timed):
        t = time.perf_counter_ns()
        if raw_copies:
            rt.cudaStreamWaitEvent(CUR, ready_h[i % 2], 0)
        else:
            cur.wait_event(ready[i % 2])
        if timed:
            t = lap('wait_ready', t)
        if no_autograd:
            e = entry[i % 2]
            e.graph.replay()
            loss = e.loss
        else:
            if lean:                 # persistent leaf per slot
                f = leaves[i % 2]
                f.grad = None
            else:
                f = slots[i % 2].detach().requires_grad_(True)
            loss = crit(f)
        if timed:
            t = lap('forward', t)
        if order_d2h_first:
            d2h(loss)
            if lean:
                loss_read.record(cur)
            if timed:
                t = lap('d2h', t)
        prefetch(i + 1)
        if timed:
            t = lap('prefetch', t)
        if not no_autograd:
            if use_ones:
                loss.backward(ones)
            else:
                loss.backward()
            if timed:
                t = lap('backward', t)
        if raw_copies:
            rt.cudaEventRecord(freed_h[i % 2], CUR)
        else:
            freed[i % 2].record(cur)
        if timed:
            t = lap('record', t)
        if not order_d2h_first:
            d2h(loss)
            if timed:
                t = lap('d2h', t)
        if lean:                     # the step ends when the loss is on the host
            loss_read.synchronize()
        elif raw_copies:
            rt.cudaStreamSynchronize(CUR)
        else:
            cur.synchronize()
        if timed:
            t = lap('sync', t)

    prefetch(0)
    if no_autograd:                      # the graphs the module captured for the two slots
        for i in range(2):
            crit(slots[i].detach().requires_grad_(True)).backward(ones)
        per_shape = next(iter(crit._graph_cache.graphs.values()))
        entry = [per_shape[slots[0].data_ptr()], per_shape[slots[1].data_ptr()]]
    for i in range(50):
        step(i, False)
    torch.cuda.synchronize()
    seg.clear()
    t0 = time.perf_counter()
    for i in range(50, 50 + STEPS):
        step(i, False)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / STEPS * 1e6
    for i in range(STEPS):
        step(i, True)
    torch.cuda.synchronize()
    print('%-44s %7.1f us/step  %5.1f M pairs/s | ' % (name, wall, B / wall) +
          '  '.join('%s %.1f' % (k, v / STEPS / 1e3) for k, v in seg.items()), flush=True)


variant('bench.py today', False, False, False)
variant('+ loss D2H issued before the backward', True, False, False)
variant('+ backward(ones)', True, True, False)
variant('+ raw cudart copies/events', True, True, True)
variant('raw copies, D2H last', False, True, True)
variant('bench.py now: persistent leaves, event wait', True, False, True, lean=True)
variant('same with backward(ones)', True, True, True, lean=True)
with torch.autograd.set_multithreading_enabled(False):      # backward runs in the calling thread: no hand-over
    variant('bench.py now + single-threaded autograd', True, False, True, lean=True)
    variant('same with backward(ones)', True, True, True, lean=True)
variant('floor: graph replay + raw copies, no autograd', True, True, True, no_autograd=True)

# bench.py's own end-to-end step function, timed the same way (wall clock) in this process
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench   # noqa: E402
for pool in (8, 64):
    _, step_e2e, _, _, _ = bench.make_loss_runner(1, 0, 'bf16', pool)
    for i in range(50):
        step_e2e(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(50, 50 + STEPS):
        step_e2e(i)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / STEPS * 1e6
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(50 + STEPS, 100 + STEPS):
        step_e2e(i)
    e1.record()
    torch.cuda.synchronize()
    print('bench.make_loss_runner step_e2e, pool %2d: %7.1f us/step wall over %d steps, %7.1f us/step by events over 50'
          % (pool, wall, STEPS, e0.elapsed_time(e1) * 1e3 / 50), flush=True)
