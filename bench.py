#!/usr/bin/env python
"""Benchmark of the two hot paths (contract: see the task statement / DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload both|ntxent|syncbn]

Prints ONE JSON line (rank 0).  Primary metric: NT-Xent fwd+bwd pairs/s at global 2N=8192, D=128
(BASELINE.json configs[3], run on N GPUs with rows sharded); the `syncbn` object carries the second
metric (SyncBN fwd+bwd HBM GB/s on [256,256,8,8] NHWC bf16 per GPU, configs[1]).

A "step" = one forward + backward of the loss on one batch of synthetic embeddings (or of the
batch-norm layer on one activation tensor).  `value` is timed with inputs resident in HBM, `e2e`
through the public module API from pinned host buffers (H2D of the inputs and D2H of the results
inside the timed region).  Timed with CUDA events on the launching stream, max over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'NT-Xent fwd+bwd pairs/s at global 2N=8192 D=128'
B_GLOBAL, N_VIEWS, DIM, TEMP = 4096, 2, 128, 0.5            # BASELINE.json configs[3]; T as main_supcon.py:214
BN_SHAPE = (256, 256, 8, 8)                                   # BASELINE.json configs[1], per GPU


def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=p['hbm_gbs'], tf=p['bf16_tflops'], tf_sus=p.get('bf16_tflops_sustained'), src='measured')
    return dict(hbm=6650.0, tf=1590.0, tf_sus=1400.0, src='fallback')


def profiled_traffic(kernel_prefix):
    """DRAM bytes per launch of a kernel from the committed `ncu --set full` summary (profiles/traffic.json,
    written by tools/make_profile_summary.py); None when no capture of that kernel is committed."""
    path = os.path.join(ROOT, 'profiles', 'traffic.json')
    if not os.path.exists(path):
        return None
    with open(path) as f:
        t = json.load(f)
    for k, v in t.items():
        if k.startswith(kernel_prefix):
            return v['dram_bytes']
    return None


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the GPU is under the bench's load."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.samples, self.proc = [], None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'note': 'nvidia-smi unavailable'}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ts, line in self.samples:
            if ts < t0 or ts > t1 + 0.15:
                continue
            parts = [p.strip() for p in line.split(',')]
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(sm)}


# ----------------------------------------------------------------------------- helpers
def dist_setup(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29533')
        torch.cuda.set_device(local)
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    else:
        torch.cuda.set_device(0)
    return world, rank, local


def max_over_ranks(ms, world):
    import torch
    import torch.distributed as dist
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device='cuda')
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def time_steps(step_fn, steps, warmup, world, graph_pool=0):
    """W untimed steps, then exactly K steps between barrier+synchronize, CUDA events, max over ranks.

    graph_pool > 0: the K steps are replays of a CUDA graph that holds g consecutive steps (g = the
    largest divisor of K not above graph_pool, each step on its own buffer set), so the launch-bound
    host side of the eager module API does not starve the GPU.  Returns (ms, how)."""
    import torch
    for i in range(warmup):
        step_fn(i)
    barrier(world)
    how, graph, reps = 'eager launches', None, 0
    if graph_pool > 0 and os.environ.get('SIMCLR_B200_BENCH_NO_GRAPH') != '1':
        g = max(d for d in range(1, min(graph_pool, steps) + 1) if steps % d == 0)
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for i in range(g):
                    step_fn(i)
            graph.replay()
            torch.cuda.synchronize()
            reps = steps // g
            how = 'CUDA graph of %d consecutive steps replayed %d times' % (g, reps)
        except Exception as e:      # say so instead of hiding it
            graph = None
            how = 'eager launches (graph capture failed: %s)' % type(e).__name__
            torch.cuda.synchronize()
        if world > 1:               # all ranks must take the same path (the steps contain collectives)
            import torch.distributed as dist
            flag = torch.tensor([1 if graph is not None else 0], device='cuda')
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 0 and graph is not None:
                graph, how = None, 'eager launches (graph capture failed on another rank)'
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if graph is not None:
        for _ in range(reps):
            graph.replay()
    else:
        for i in range(steps):
            step_fn(warmup + i)
    e1.record()
    barrier(world)
    return max_over_ranks(e0.elapsed_time(e1), world), how


# ----------------------------------------------------------------------------- NT-Xent (ours)
def make_loss_runner(world, rank, engine, pool, b_local=None):
    """Returns step functions for the loss on this rank's shard of the 2N=8192 global batch (b_local given: that
    many pairs per rank instead -- the weak-scaling series, 2N = 2 * b_local * world)."""
    import torch
    import torch.nn.functional as F
    from simclr_pytorch_distributed_b200 import SupConLoss
    b_local = B_GLOBAL // world if b_local is None else b_local
    dtype = torch.bfloat16 if engine == 'bf16' else torch.float32
    g = torch.Generator().manual_seed(1234 + rank)
    host = [F.normalize(torch.randn(b_local, N_VIEWS, DIM, generator=g), dim=-1).to(dtype).pin_memory()
            for _ in range(pool)]
    dev = [h.cuda().requires_grad_(True) for h in host]
    if world == 1:
        crit = SupConLoss(temperature=TEMP, engine=engine)
        fwd = lambda f: crit(f)                                                     # noqa: E731
        # end to end: the module's graph mode (one CUDA-graph launch per step instead of four C-ABI calls + autograd)
        crit_g = SupConLoss(temperature=TEMP, engine=engine, graph=True)
        fwd_e2e = crit_g
    else:
        from simclr_pytorch_distributed_b200.gather import DistributedSupConLoss
        crit = DistributedSupConLoss(temperature=TEMP, engine=engine, normalize=False)
        fwd = lambda f: crit(f)                                                     # noqa: E731
        crit_g = DistributedSupConLoss(temperature=TEMP, engine=engine, normalize=False, graph=True)
        fwd_e2e = crit_g

    ones = torch.ones((), dtype=dtype, device='cuda')       # upstream gradient of the scalar loss, allocated once

    def step_resident(i):
        f = dev[i % pool]
        f.grad = None
        fwd(f).backward(ones)

    # End to end: every step copies ITS embeddings from pinned host memory and the host reads the loss.  The
    # copy of step i+1 is issued on a side stream before step i computes (what a training loop's prefetcher
    # does), so it overlaps the kernels; the gradient stays on the device, where the encoder's backward
    # consumes it (reading the 2 MiB gradient back every step would only measure PCIe).  The loss is copied to
    # the host as soon as the forward has been issued and the step ends when that copy has landed AND the
    # backward has been issued (tools/prof_host_loss.py: where the host time of this step goes).
    loss_host = torch.empty((), dtype=dtype).pin_memory()     # the loss has the features' dtype (losses.py:91)
    copy_stream = torch.cuda.Stream()
    slots = [torch.empty(b_local, N_VIEWS, DIM, dtype=dtype, device='cuda') for _ in range(2)]
    leaves = [s.detach().requires_grad_(True) for s in slots]
    ready = [torch.cuda.Event() for _ in range(2)]          # H2D of the slot has landed
    freed = [torch.cuda.Event() for _ in range(2)]          # the step that used the slot is done with it
    loss_read = torch.cuda.Event()                          # the loss of the step is on the host
    state = {'primed': -1}
    for e in freed + ready + [loss_read]:
        e.record()
    nbytes = b_local * N_VIEWS * DIM * host[0].element_size()
    # The copies and event calls of the prefetcher go through cudart directly where cuda-python is importable:
    # torch's stream context manager + copy_ cost ~20 us of host time per step, these ~4 us.
    try:
        if os.environ.get('SIMCLR_B200_BENCH_TORCH_COPIES') == '1':
            raise ImportError
        from cuda.bindings import runtime as rt
        h2d, d2h_kind = rt.cudaMemcpyKind.cudaMemcpyHostToDevice, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost
        cs = copy_stream.cuda_stream
        ready_h, freed_h = [e.cuda_event for e in ready], [e.cuda_event for e in freed]
        host_ptr, slot_ptr, loss_host_ptr = [h.data_ptr() for h in host], [t.data_ptr() for t in slots], loss_host.data_ptr()
    except (ImportError, AttributeError):       # no cuda-python (or another layout of it): torch's own calls
        rt = None

    def prefetch(j):
        s = j % 2
        if rt is not None:
            rt.cudaStreamWaitEvent(cs, freed_h[s], 0)
            err, = rt.cudaMemcpyAsync(slot_ptr[s], host_ptr[j % pool], nbytes, h2d, cs)
            rt.cudaEventRecord(ready_h[s], cs)
            if int(err) != 0:
                raise RuntimeError('cudaMemcpyAsync (prefetch): %s' % err)
        else:
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(freed[s])
                slots[s].copy_(host[j % pool], non_blocking=True)
                ready[s].record(copy_stream)
        state['primed'] = j

    cur = torch.cuda.current_stream()                         # the bench never changes its stream
    cur_h, loss_bytes = cur.cuda_stream, loss_host.element_size()

    def step_e2e(i):
        if state['primed'] != i:
            prefetch(i)                                       # first step of a run: nothing was prefetched
        s = i % 2
        f = leaves[s]
        f.grad = None
        if rt is not None:
            rt.cudaStreamWaitEvent(cur_h, ready_h[s], 0)
            loss = fwd_e2e(f)
            err, = rt.cudaMemcpyAsync(loss_host_ptr, loss.data_ptr(), loss_bytes, d2h_kind, cur_h)
            if int(err) != 0:
                raise RuntimeError('cudaMemcpyAsync (loss): %s' % err)
        else:
            cur.wait_event(ready[s])
            loss = fwd_e2e(f)
            loss_host.copy_(loss.detach(), non_blocking=True)
        loss_read.record(cur)
        prefetch(i + 1)                                       # issued while the GPU works on step i
        loss.backward()
        if rt is not None:
            rt.cudaEventRecord(freed_h[s], cur_h)
        else:
            freed[s].record(cur)
        loss_read.synchronize()                               # the caller reads the loss every step

    return step_resident, step_e2e, nbytes, loss_host.element_size(), b_local


def loss_kernel_breakdown(engine, steps):
    """CUDA-event times of the launches of one fwd+bwd through the C ABI (N=1 shapes, the product path: bf16
    features used in place), for the roofline.  The backward's main kernel is timed alone with two events the
    library records around it (scb_debug_set_events)."""
    import torch
    import torch.nn.functional as F
    from simclr_pytorch_distributed_b200 import losses as L, _lib
    lib = _lib.load()
    dev = torch.device('cuda')
    dtype = torch.bfloat16 if engine == 'bf16' else torch.float32
    z = F.normalize(torch.randn(B_GLOBAL, N_VIEWS, DIM, device=dev), dim=-1).to(dtype)
    plan = L.LossPlan(B_GLOBAL, N_VIEWS, DIM, engine=L._ENGINES[engine])
    plan.scalar_dtype = _lib.dtype_code(dtype)
    tc = engine != 'fp32'
    if tc:
        plan.row_layout = _lib.ROWS_SAMPLE_MAJOR
        k = z.view(plan.n_k, plan.ldk)
    else:
        k = torch.empty((plan.n_k, plan.ldk), dtype=plan.k_torch, device=dev)
        L.launch_prep(lib, z, plan, False, k, None, None)
    stats = torch.empty((4, plan.n_k), dtype=torch.float32, device=dev)
    loss = torch.empty(2, dtype=torch.float32, device=dev)
    gout = torch.ones((), dtype=dtype, device=dev)
    dk = torch.empty((plan.rows_local, plan.ldk), dtype=torch.float32, device=dev)
    dz = torch.empty_like(z)
    prob = L._problem(plan, k, None, None, None, TEMP, 0.07)
    names = ['forward(+finalize)', 'backward(+reduce/finish)', 'backward main kernel']
    tot = [0.0] * 3
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    em = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for e in em:
        e.record()                      # creates the underlying cudaEvent_t
    torch.cuda.synchronize()
    if tc:
        lib.scb_debug_set_events(em[0].cuda_event, em[1].cuda_event)
    try:
        for it in range(steps + 3):
            torch.cuda._sleep(400000)      # let the host run ahead so the events see back-to-back kernels
            ev[0].record()
            L.launch_forward(lib, prob, plan, stats[0], stats[1], stats[2], loss, dev)
            ev[1].record()
            if tc:
                L.launch_backward_features(lib, prob, plan, stats[0], stats[2], gout, z, False, None, dz, dev)
            else:
                L.launch_backward(lib, prob, plan, stats[0], stats[2], gout, dk, dev)
                L.launch_finish(lib, dk, z, plan, False, None, dz)
            ev[2].record()
            torch.cuda.synchronize()
            if it >= 3:
                tot[0] += ev[0].elapsed_time(ev[1])
                tot[1] += ev[1].elapsed_time(ev[2])
                tot[2] += em[0].elapsed_time(em[1]) if tc else ev[1].elapsed_time(ev[2])
    finally:
        lib.scb_debug_set_events(None, None)
    return {n: t / steps * 1e3 for n, t in zip(names, tot)}     # microseconds


# ----------------------------------------------------------------------------- SyncBN (ours)
def make_bn_runner(world, rank, pool):
    import torch
    from simclr_pytorch_distributed_b200 import SyncBatchNorm
    N, C, H, W_ = BN_SHAPE
    g = torch.Generator().manual_seed(4321 + rank)
    mu = (torch.rand(1, C, 1, 1, generator=g) * 4 - 2)
    sg = (torch.rand(1, C, 1, 1, generator=g) * 1.5 + 0.5)
    mk = lambda: (torch.randn(N, C, H, W_, generator=g) * sg + mu).to(torch.bfloat16).contiguous(   # noqa: E731
        memory_format=torch.channels_last)
    host_x = [mk().pin_memory() for _ in range(pool)]
    host_dy = [torch.randn(N, C, H, W_, generator=g).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last).pin_memory() for _ in range(pool)]
    dev_x = [h.cuda().requires_grad_(True) for h in host_x]
    dev_dy = [h.cuda() for h in host_dy]
    bn = SyncBatchNorm(C).cuda()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.normal_(0, 0.1)

    def step_resident(i):
        x = dev_x[i % pool]
        x.grad = None
        bn.weight.grad = None                   # optimizer.zero_grad() of the training loop (main_supcon.py:322):
        bn.bias.grad = None                     # without it autograd adds two accumulation kernels per step
        bn(x).backward(dev_dy[i % pool])

    y_host = torch.empty_like(host_x[0]).pin_memory()
    dx_host = torch.empty_like(host_x[0]).pin_memory()

    def step_e2e(i):
        x = host_x[i % pool].cuda(non_blocking=True).requires_grad_(True)
        dy = host_dy[i % pool].cuda(non_blocking=True)
        bn.weight.grad = None
        bn.bias.grad = None
        y = bn(x)
        y.backward(dy)
        y_host.copy_(y.detach(), non_blocking=True)
        dx_host.copy_(x.grad, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    S = N * C * H * W_ * 2
    return step_resident, step_e2e, S, bn


def bn_kernel_breakdown(steps, pool=8):
    import torch
    from simclr_pytorch_distributed_b200.sync_batchnorm import _Kernels as K, _layout
    N, C, H, W_ = BN_SHAPE
    dev = torch.device('cuda')
    xs = [(torch.randn(N, C, H, W_, device=dev) * 1.5 + 0.5).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last) for _ in range(pool)]
    dys = [torch.randn(N, C, H, W_, device=dev).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last) for _ in range(pool)]
    ys = [torch.empty_like(x) for x in xs]
    dxs = [torch.empty_like(x) for x in xs]
    w = torch.rand(C, device=dev) + 0.5
    b = torch.randn(C, device=dev) * 0.1
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    st = torch.empty(2 * C + 1, device=dev)
    saved = torch.empty(2 * C + 1, device=dev)
    out = torch.empty(4 * C, device=dev)
    _, geom = _layout(xs[0])
    fused = K.fused_supported(xs[0], geom, 1)
    names = ['forward_fused', 'backward_fused'] if fused else ['stats', 'apply', 'backward_reduce', 'backward_apply']
    tot = [0.0] * len(names)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    for it in range(steps + 3):
        x, dy, y, dx = xs[it % pool], dys[it % pool], ys[it % pool], dxs[it % pool]
        torch.cuda._sleep(400000)      # let the host run ahead so the events see back-to-back kernels
        ev[0].record()
        if fused:
            K.forward_fused(x, y, geom, w, b, rm, rv, 0.1, 1e-5, saved, 0)
            ev[1].record()
            K.backward_fused(dy, x, dx, geom, saved, w, b, out[:2 * C], out[2 * C:3 * C], out[3 * C:], 0)
            ev[2].record()
        else:
            K.stats(x, geom, st)
            ev[1].record()
            K.apply(x, y, geom, st, 1, w, b, rm, rv, 0.1, 1e-5, saved, 0)
            ev[2].record()
            K.backward_reduce(dy, x, geom, saved, w, b, out[:2 * C], out[2 * C:3 * C], out[3 * C:], 0)
            ev[3].record()
            K.backward_apply(dy, x, dx, geom, saved, w, b, out[:2 * C], 0)
            ev[4].record()
        torch.cuda.synchronize()
        if it >= 3:
            for j in range(len(names)):
                tot[j] += ev[j].elapsed_time(ev[j + 1])
    return {n: t / steps * 1e3 for n, t in zip(names, tot)}


def loss_fp32_variant(args, pk):
    import torch
    from simclr_pytorch_distributed_b200 import SupConLoss, _lib
    lib = _lib.load()
    pool = 32
    g = torch.Generator().manual_seed(4242)
    zs = [(torch.randn(B_GLOBAL, N_VIEWS, DIM, generator=g) * 2.0 + 0.3).cuda().requires_grad_(True) for _ in range(pool)]
    crit = SupConLoss(temperature=TEMP, normalize=True)          # engine 'auto'
    ones = torch.ones((), device='cuda')

    def step(i):
        zs[i % pool].grad = None
        crit(zs[i % pool]).backward(ones)
    step(0)
    c0 = lib.scb_launch_count()
    step(1)
    launches = lib.scb_launch_count() - c0
    ms, how = time_steps(step, args.steps, args.warmup, 1, graph_pool=pool)
    per = ms / args.steps
    flops = 6.0 * (N_VIEWS * B_GLOBAL) ** 2 * DIM
    ach = flops / (per * 1e-3) / 1e12
    # parity of exactly this call against the fp64 dense algorithm on the raw fp32 rows
    from oracle.dense_port import dense_supcon_loss
    import torch.nn.functional as F
    f = zs[0].detach().clone().requires_grad_(True)
    loss = crit(f)
    loss.backward()
    r = zs[0].detach().double().requires_grad_(True)
    ref = dense_supcon_loss(F.normalize(r, dim=-1), temperature=TEMP)
    ref.backward()
    return {'metric': METRIC, 'value': B_GLOBAL / (per * 1e-3), 'unit': 'pairs/s', 'ms_per_step': per,
            'dtype': 'bf16x3 (bf16 hi+lo operands, fp32 accumulate)', 'gpu_launches': int(launches),
            'config': {'workload': 'raw fp32 embeddings [4096,2,128], normalize=True (fused F.normalize and its '
                                   'backward), engine auto', 'launch': how},
            'roofline': {'bound': 'tensor', 'achieved': ach, 'peak': pk['tf'] / 3.0, 'unit': 'TFLOP/s',
                         'frac': ach / (pk['tf'] / 3.0), 'peak_source': pk['src'] + ' burst bf16 / 3 (three MMAs per '
                         'algorithmic product)'},
            'parity': {'loss_rel_err': abs(float(loss) - float(ref)) / abs(float(ref)),
                       'grad_rel_fro_err': float((f.grad.double() - r.grad).norm() / r.grad.norm()), 'tolerance': 1e-4}}


# ----------------------------------------------------------------------------- parity of what is timed
def all_gather_cat(t, world):
    import torch
    import torch.distributed as dist
    if world == 1:
        return t
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t.contiguous())
    return torch.cat(parts, 0)


def loss_parity(world, rank, engine):
    """Loss and this rank's gradient on the bench's own inputs against the reference's dense algorithm
    (oracle/dense_port.py: /root/reference/losses.py:52-91 restated, eager PyTorch, fp64, on the GPU) fed the
    all-gathered embeddings.  Checker only: nothing of this is timed as ours."""
    import torch
    import torch.nn.functional as F
    from oracle.dense_port import dense_supcon_loss
    from simclr_pytorch_distributed_b200 import SupConLoss
    b_local = B_GLOBAL // world
    dtype = torch.bfloat16 if engine == 'bf16' else torch.float32
    g = torch.Generator().manual_seed(99 + rank)
    z = F.normalize(torch.randn(b_local, N_VIEWS, DIM, generator=g), dim=-1).to(dtype).cuda()
    f = z.clone().requires_grad_(True)
    if world == 1:
        loss = SupConLoss(temperature=TEMP, engine=engine)(f)
    else:
        from simclr_pytorch_distributed_b200.gather import DistributedSupConLoss
        loss = DistributedSupConLoss(temperature=TEMP, engine=engine, normalize=False)(f)
    loss.backward()
    z_all = all_gather_cat(z, world)
    out = None
    if rank == 0:
        ref_in = z_all.double().requires_grad_(True)
        ref = dense_supcon_loss(ref_in, temperature=TEMP)
        ref.backward()
        gref = ref_in.grad[:b_local]
        out = {'loss_rel_err': abs(float(loss) - float(ref)) / abs(float(ref)),
               'grad_rel_fro_err': float((f.grad.double() - gref).norm() / gref.norm()),
               'tolerance': 2e-3 if engine == 'bf16' else 1e-4,
               'against': 'dense fp64 algorithm of losses.py:52-91 (oracle/dense_port.py, eager torch on the GPU) on the '
                          'all-gathered bf16 embeddings; gradient of rank 0 vs its slice'}
        del ref_in, ref
    torch.cuda.empty_cache()
    return out


def bn_parity(world, rank):
    """SyncBatchNorm y / dx / running statistics on per-rank tensors against stock torch.nn.SyncBatchNorm (NCCL,
    world > 1) or torch.nn.BatchNorm2d (world 1): bf16 NHWC cfg 2 (within one bf16 ulp) and fp32 (1e-5)."""
    import torch
    from simclr_pytorch_distributed_b200 import SyncBatchNorm
    res = {}
    for name, dtype, shape in (('bf16_cfg2', torch.bfloat16, BN_SHAPE), ('fp32', torch.float32, (64, 256, 8, 8))):
        N, C, H, W_ = shape
        g = torch.Generator().manual_seed(777 + rank)
        x = (torch.randn(shape, generator=g) * 1.5 + 0.5).to(dtype).cuda().contiguous(memory_format=torch.channels_last)
        dy = torch.randn(shape, generator=g).to(dtype).cuda().contiguous(memory_format=torch.channels_last)
        ours = SyncBatchNorm(C).cuda()
        ref = (torch.nn.SyncBatchNorm(C) if world > 1 else torch.nn.BatchNorm2d(C)).cuda()
        with torch.no_grad():
            ours.weight.uniform_(0.5, 1.5)
            ours.bias.normal_(0, 0.1)
        if world > 1:
            import torch.distributed as dist
            for t in (ours.weight, ours.bias):
                dist.broadcast(t.data, 0)
        ref.load_state_dict(ours.state_dict())
        xo, xr = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
        yo = ours(xo)
        yo.backward(dy)
        yr = ref(xr)
        yr.backward(dy)
        ulp = 2.0 ** -7 if dtype == torch.bfloat16 else 0.0

        def err(a, b):      # worst error beyond one output ulp, relative to the largest reference magnitude
            a, b = a.double(), b.double()
            return float(((a - b).abs() - ulp * b.abs()).clamp_min(0).max() / b.abs().max())
        res[name] = {'y_err': err(yo, yr), 'dx_err': err(xo.grad, xr.grad),
                     'running_mean_err': float((ours.running_mean - ref.running_mean).abs().max()),
                     'running_var_rel_err': float(((ours.running_var - ref.running_var).abs() / ref.running_var).max()),
                     'dgamma_rel_err': float((ours.weight.grad - ref.weight.grad).norm() / ref.weight.grad.norm())}
    res['tolerance'] = 1e-5
    res['against'] = ('torch.nn.SyncBatchNorm over NCCL' if world > 1 else 'torch.nn.BatchNorm2d') + \
                     ' on the same per-rank tensors; *_err = error beyond one ulp of the output dtype / max |reference|'
    return res


# ----------------------------------------------------------------------------- on-box GPU bar
def gpu_loss_baseline(steps=10):
    """The reference's dense algorithm in eager PyTorch on THIS GPU (fp32, as main_supcon.py runs it; and with TF32
    matmuls allowed), full 2N = 8192 problem on one GPU -- what every rank of the reference computes."""
    import torch
    import torch.nn.functional as F
    from oracle.dense_port import dense_supcon_fwd_bwd
    z = F.normalize(torch.randn(B_GLOBAL, N_VIEWS, DIM, device='cuda'), dim=-1)
    out = {}
    for name, tf32 in (('fp32', False), ('tf32', True)):
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        try:
            for _ in range(3):
                dense_supcon_fwd_bwd(z, temperature=TEMP)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                dense_supcon_fwd_bwd(z, temperature=TEMP)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            out[name] = {'ms_per_step': ms, 'value': B_GLOBAL / (ms * 1e-3), 'unit': 'pairs/s'}
        finally:
            torch.backends.cuda.matmul.allow_tf32 = old
    out['what'] = 'oracle/dense_port.py (losses.py:52-91 restated) in eager PyTorch on the B200, 2N=8192 on one GPU'
    torch.cuda.empty_cache()
    return out


def gpu_bn_baseline(world, steps=20, pool=12):
    """Stock torch.nn.SyncBatchNorm (NCCL, world > 1) or the four ATen ops it runs (world 1) on cfg 2."""
    import torch
    N, C, H, W_ = BN_SHAPE
    xs = [(torch.randn(BN_SHAPE, device='cuda') * 1.5 + 0.5).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last).requires_grad_(True) for _ in range(pool)]
    dys = [torch.randn(BN_SHAPE, device='cuda').to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
           for _ in range(pool)]
    if world > 1:
        bn = torch.nn.SyncBatchNorm(C).cuda()

        def step(i):
            xs[i % pool].grad = None
            bn(xs[i % pool]).backward(dys[i % pool])
        what = 'torch.nn.SyncBatchNorm fwd+bwd over NCCL, eager'
    else:
        w, b = torch.ones(C, device='cuda'), torch.zeros(C, device='cuda')
        count = torch.full((1,), N * H * W_, dtype=torch.int32, device='cuda')

        def step(i):
            x, dy = xs[i % pool].detach(), dys[i % pool]
            mean, invstd = torch.batch_norm_stats(x, 1e-5)
            torch.batch_norm_elemt(x, w, b, mean, invstd, 1e-5)
            sd, sx, _, _ = torch.batch_norm_backward_reduce(dy, x, mean, invstd, w, True, True, True)
            torch.batch_norm_backward_elemt(dy, x, mean, invstd, w, sd, sx, count)
        what = 'the four ATen ops of torch/nn/modules/_functions.py:39,122,144,170, eager'
    for i in range(5):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    S = N * C * H * W_ * 2
    return {'ms_per_step': ms, 'value': 8.0 * S / (ms * 1e-3) / 1e9, 'unit': 'GB/s per GPU', 'what': what}


def head_tail_bench(iters=20):
    """the projection-head tail at the cfg-3 size (512 samples x 2 views per GPU, 2048 -> 128, bf16),
    this library's one kernel against the torch ops it replaces (Linear, normalize, view stacking), both as CUDA
    graphs of `iters` forward calls; the step that follows (loss on the tail's output, used in place) is not part of
    it."""
    import torch
    import torch.nn.functional as F
    from simclr_pytorch_distributed_b200 import ProjectionTail
    M, K, N = 1024, 2048, 128
    hs = [torch.relu(torch.randn(M, K, device='cuda')).to(torch.bfloat16) for _ in range(8)]
    lin = torch.nn.Linear(K, N).cuda().to(torch.bfloat16)
    tail = ProjectionTail.from_linear(lin, n_views=2)

    def ours(i):
        return tail(hs[i % 8])

    def stock(i):
        y = F.normalize(lin(hs[i % 8]), dim=1)
        return torch.cat([y[:M // 2].unsqueeze(1), y[M // 2:].unsqueeze(1)], dim=1)      # main_supcon.py:285-286

    out = {}
    with torch.no_grad():
        for name, fn in (('ours_us', ours), ('torch_us', stock)):
            for i in range(3):
                fn(i)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for i in range(iters):
                    fn(i)
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            out[name] = e0.elapsed_time(e1) / (5 * iters) * 1e3
        err = float((ours(0).float() - stock(0).float()).abs().max())
    out['max_abs_diff_vs_torch_bf16'] = err
    out['what'] = ('normalize(Linear(h)) + view stacking, h [1024, 2048] bf16 -> [512, 2, 128] bf16: one tcgen05 kernel '
                   '(split-K over a cluster of 8, DSMEM reduction, norm in the epilogue) vs cuBLAS + 4 torch kernels')
    return out



# ----------------------------------------------------------------------------- configs[2]: the model-level step
def run_model(args, world, rank):
    """BASELINE configs[2]: SimCLR ResNet-50 + MLP head (tests/carrier_resnet.py restates networks/resnet_big.py),
    CIFAR-shape synthetic crops, per-GPU batch 256 (two crops -> 512 images), bf16 channels_last, DDP when N > 1.
    One step = forward, loss, backward, SGD update -- /root/reference/main_supcon.py:266-325.  Three arms on the same
    data: this library swapped in as the reference does it (`convert_sync_batchnorm` + the loss module), the same with
    ReLU / add+ReLU folded into the batch-norm kernels, and the stock path (torch.nn.SyncBatchNorm / BatchNorm2d +
    all-gather + the dense loss on every rank)."""
    import copy
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    from carrier_resnet import SupConResNet50
    from oracle.dense_port import dense_supcon_loss
    from simclr_pytorch_distributed_b200 import SupConLoss, SyncBatchNorm
    from simclr_pytorch_distributed_b200.gather import DistributedSupConLoss
    bsz = 256
    dev = torch.device('cuda')
    torch.manual_seed(0)
    base = SupConResNet50().to(dev).to(memory_format=torch.channels_last).to(torch.bfloat16)
    g = torch.Generator().manual_seed(2024 + rank)
    pool = [torch.randn(2 * bsz, 3, 32, 32, generator=g).to(torch.bfloat16).cuda().contiguous(
        memory_format=torch.channels_last) for _ in range(4)]

    def wrap(m):
        return torch.nn.parallel.DistributedDataParallel(m, device_ids=[dev.index]) if world > 1 else m

    def make(arm):
        if arm == 'stock':
            m = copy.deepcopy(base)
            if world > 1:
                m = torch.nn.SyncBatchNorm.convert_sync_batchnorm(m)

            def crit(feats):                       # main_supcon.py:268-290: gather, re-insert the local slice, dense loss
                if world > 1:
                    parts = [torch.empty_like(feats) for _ in range(world)]
                    dist.all_gather(parts, feats.detach())
                    parts[rank] = feats
                    f1 = torch.cat([p[:bsz] for p in parts])
                    f2 = torch.cat([p[bsz:] for p in parts])
                else:
                    f1, f2 = feats[:bsz], feats[bsz:]
                return dense_supcon_loss(torch.cat([f1.unsqueeze(1), f2.unsqueeze(1)], 1).float(), temperature=TEMP)
            normalize_in_model = True
        else:
            if arm == 'ours_fused':
                # ReLU / add+ReLU inside the batch-norm kernels, and the end of the head (Linear, normalise, view
                # stacking) as one kernel whose bf16 output the loss uses in place
                m = SupConResNet50(bn=SyncBatchNorm, fuse=True, tail=True).to(dev).to(
                    memory_format=torch.channels_last).to(torch.bfloat16)
                m.load_state_dict(base.state_dict())
                loss_mod = (DistributedSupConLoss(temperature=TEMP, normalize=False) if world > 1
                            else SupConLoss(temperature=TEMP))

                def crit(feats):
                    return loss_mod(feats)
            else:
                m = SyncBatchNorm.convert_sync_batchnorm(copy.deepcopy(base))
                loss_mod = (DistributedSupConLoss(temperature=TEMP, normalize=True) if world > 1
                            else SupConLoss(temperature=TEMP, normalize=True))

                def crit(feats):                       # the fused normalise replaces main_supcon.py:283
                    return loss_mod(torch.stack([feats[:bsz], feats[bsz:]], 1))
            normalize_in_model = False
        m = wrap(m)
        opt = torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)

        def step(i):
            feats = m(pool[i % len(pool)], normalize=normalize_in_model)
            loss = crit(feats)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
            return loss
        return step

    def bn_share(step):
        try:
            from torch.profiler import profile, ProfilerActivity
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                for i in range(2):
                    step(i)
                torch.cuda.synchronize()
            tot = bn = 0.0
            for ev in prof.key_averages():
                t = getattr(ev, 'device_time_total', None) or getattr(ev, 'cuda_time_total', 0.0)
                tot += t
                if 'bn_' in ev.key or 'batch_norm' in ev.key:
                    bn += t
            return (bn / tot if tot > 0 else None), tot / 2 / 1e3        # share, GPU-busy ms per step (sum of kernel times)
        except Exception as e:      # say so
            return 'profiler unavailable: %s' % type(e).__name__, None

    res = {}
    for arm in ('stock', 'ours_swap', 'ours_fused'):
        step = make(arm)
        ms, _ = time_steps(step, min(args.steps, 20), max(3, min(args.warmup, 5)), world)
        per = ms / min(args.steps, 20)
        share, busy = bn_share(step)
        res[arm] = {'ms_per_step': per, 'images_per_s': 2 * bsz * world / (per * 1e-3),
                    'batch_norm_share_of_gpu_time': share, 'gpu_busy_ms_per_step': busy,
                    'loss_after': float(step(0).detach())}
        del step
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps({
            'metric': 'SimCLR ResNet-50 training step, images/s', 'unit': 'images/s', 'n_gpus': world,
            'value': res['ours_fused']['images_per_s'], 'ms_per_step': res['ours_fused']['ms_per_step'],
            'higher_is_better': True, 'scaling': 'weak', 'data': 'synthetic', 'dtype': 'bf16', 'vs_baseline': None,
            'config': {'workload': 'SimCLR ResNet-50 + MLP head, 32x32 crops, per-GPU batch 256 (512 images), bf16 '
                                   'channels_last, fwd + loss + bwd + SGD' + (', DDP' if world > 1 else ''),
                       'launch': 'eager'},
            'arms': res,
            'speedup_over_stock': {k: res['stock']['ms_per_step'] / res[k]['ms_per_step'] for k in ('ours_swap', 'ours_fused')},
        }), flush=True)


# ----------------------------------------------------------------------------- CPU baselines
def cpu_loss_baseline(budget_s=15.0, max_steps=20):
    """The reference's dense CPU algorithm (oracle/dense_port.py, a port: /root/reference does not
    travel to the GPU box) on the same [4096, 2, 128] fp32 workload, all host threads."""
    import torch
    import torch.nn.functional as F
    from oracle.dense_port import dense_supcon_fwd_bwd
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    z = F.normalize(torch.randn(B_GLOBAL, N_VIEWS, DIM, generator=torch.Generator().manual_seed(1234)), dim=-1)
    dense_supcon_fwd_bwd(z, temperature=TEMP)          # warm-up
    times, t_start = [], time.time()
    while len(times) < max_steps and (time.time() - t_start < budget_s or len(times) < 2):
        t0 = time.perf_counter()
        dense_supcon_fwd_bwd(z, temperature=TEMP)
        times.append(time.perf_counter() - t0)
    times.sort()
    med = times[len(times) // 2]
    return med, len(times), cores


def cpu_bn_baseline(budget_s=8.0, max_steps=50):
    import torch
    from oracle.dense_port import bn_train_fwd_bwd
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    N, C, H, W_ = BN_SHAPE
    x = (torch.randn(N, C, H, W_) * 1.5 + 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(N, C, H, W_).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w, b = torch.ones(C), torch.zeros(C)
    rm, rv = torch.zeros(C), torch.ones(C)
    bn_train_fwd_bwd(x, dy, w, b, rm, rv)
    times, t_start = [], time.time()
    while len(times) < max_steps and (time.time() - t_start < budget_s or len(times) < 2):
        t0 = time.perf_counter()
        bn_train_fwd_bwd(x, dy, w, b, rm, rv)
        times.append(time.perf_counter() - t0)
    times.sort()
    return times[len(times) // 2], len(times), cores


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    """`--impl reference`: the reference's CPU path on this box's host cores (port, see above)."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    steps, warmup = args.steps, args.warmup
    import torch
    import torch.nn.functional as F
    from oracle.dense_port import dense_supcon_fwd_bwd
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    z = F.normalize(torch.randn(B_GLOBAL, N_VIEWS, DIM, generator=torch.Generator().manual_seed(1234)), dim=-1)
    t0 = time.perf_counter()
    dense_supcon_fwd_bwd(z, temperature=TEMP)
    first = time.perf_counter() - t0
    # keep the whole run within a few minutes: cap the number of executed steps, not the workload
    budget = 150.0
    w_run = max(1, min(warmup, int(0.2 * budget / max(first, 1e-3))))
    k_run = max(2, min(steps, int(0.8 * budget / max(first, 1e-3))))
    for _ in range(w_run - 1):
        dense_supcon_fwd_bwd(z, temperature=TEMP)
    t0 = time.perf_counter()
    for _ in range(k_run):
        dense_supcon_fwd_bwd(z, temperature=TEMP)
    per = (time.perf_counter() - t0) / k_run
    val = B_GLOBAL / per
    bn_t, bn_n, _ = cpu_bn_baseline(budget_s=10.0)
    S = BN_SHAPE[0] * BN_SHAPE[1] * BN_SHAPE[2] * BN_SHAPE[3] * 2
    sample = ('full workload [4096,2,128] fp32 fwd+bwd per step; %d of the requested %d timed steps '
              'executed (after %d warm-ups)' % (k_run, steps, w_run))
    out = {
        'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': 'pairs/s', 'n_gpus': args.gpus,
        'steps': steps, 'warmup': warmup, 'ms_per_step': per * 1e3, 'higher_is_better': True,
        'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': 'NT-Xent SimCLR loss fwd+bwd, global 2N=8192 (B=4096 pairs, 2 views), D=128, '
                               'T=0.5, base_T=0.07', 'device': 'host CPU'},
        'cpu_baseline': {'value': val, 'unit': 'pairs/s', 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': val, 'unit': 'pairs/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
        'syncbn': {'metric': 'SyncBN fwd+bwd HBM GB/s', 'value': 8 * S / bn_t / 1e9, 'unit': 'GB/s',
                   'ms_per_step': bn_t * 1e3,
                   'config': {'workload': 'F.batch_norm train fwd+bwd [256,256,8,8] NHWC bf16 on host CPU'},
                   'cpu_baseline': {'value': 8 * S / bn_t / 1e9, 'unit': 'GB/s', 'cores': cores, 'kind': 'port',
                                    'sample': '%d fwd+bwd steps' % bn_n}},
    }
    print(json.dumps(out), flush=True)


# ----------------------------------------------------------------------------- main (ours)
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='both', choices=['both', 'ntxent', 'syncbn', 'syncbn-census', 'model'])
    ap.add_argument('--engine', default=os.environ.get('SIMCLR_B200_BENCH_ENGINE', 'bf16'))
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-gpu-baseline', action='store_true')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == 'reference':
        run_reference(args)
        return

    import torch
    from simclr_pytorch_distributed_b200 import _lib
    world, rank, local = dist_setup(args)
    lib = _lib.load()
    pk = peaks()
    engine = args.engine
    if engine != 'fp32':
        try:     # the tensor-core engine may not exist yet in an early round: say so, do not hide it
            from simclr_pytorch_distributed_b200 import SupConLoss
            SupConLoss(engine=engine)(torch.nn.functional.normalize(torch.randn(8, 2, 128, device='cuda'), dim=-1))
        except RuntimeError as e:
            if 'not built yet' not in str(e):
                raise
            engine = 'fp32'
    sampler = ClockSampler(local) if rank == 0 else None
    t_load0 = time.time()
    out = {}

    if args.workload == 'model':
        run_model(args, world, rank)
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()
        return
    # ------------------------------------------------ SyncBN over the layer shapes of the reference ResNet-50
    if args.workload == 'syncbn-census':
        import importlib.util
        spec = importlib.util.spec_from_file_location('bn_census', os.path.join(ROOT, 'tools', 'bn_census.py'))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        rows = mod.census(iters=max(5, min(args.steps, 20)))
        for r in rows:
            r['frac_of_hbm_peak'] = r['ours_GBs'] / pk['hbm']
            r['ours_over_aten'] = r['aten_us'] / (r['ours_graph_us'] or r['ours_us'])
        if rank == 0:
            print(json.dumps({'metric': 'SyncBN fwd+bwd HBM GB/s (8*S algorithmic bytes) per layer shape', 'unit': 'GB/s',
                              'n_gpus': world, 'higher_is_better': True, 'data': 'synthetic', 'dtype': 'bf16 I/O, f32 statistics',
                              'config': {'workload': 'SURVEY Appendix B shapes at N=512 per GPU, NHWC bf16, plus cfg 2; '
                                                     'beside the four ATen ops of stock SyncBatchNorm on the same tensors'},
                              'peak': pk['hbm'], 'peak_source': pk['src'], 'census': rows}), flush=True)
        return

    # ------------------------------------------------ NT-Xent
    if args.workload in ('both', 'ntxent'):
        pool = 64                                    # 64 x 2 MiB (bf16) of embeddings + grads > L2
        step_res, step_e2e, h2d, d2h, b_local = make_loss_runner(world, rank, engine, pool)
        c0 = lib.scb_launch_count()
        step_res(0)
        c0 = lib.scb_launch_count()
        step_res(1)
        launches = lib.scb_launch_count() - c0
        ms, how = time_steps(step_res, args.steps, args.warmup, world, graph_pool=pool)
        # one device per process: autograd may run the backward in the calling thread instead of handing it to its
        # device thread (a PyTorch switch any training script can set; 55 -> 39 us of host time per backward)
        with torch.autograd.set_multithreading_enabled(False):
            ms_e2e, _ = time_steps(step_e2e, args.steps, args.warmup, world)
        per = ms / args.steps
        A = C = N_VIEWS * B_GLOBAL
        flops_step = 6.0 * A * C * DIM
        out.update({
            'metric': METRIC, 'value': B_GLOBAL / (per * 1e-3), 'unit': 'pairs/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': per, 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None,
            'dtype': {'bf16': 'bf16', 'bf16x3': 'bf16x3 (split bf16, fp32 accumulate)', 'fp32': 'f32'}[engine],
            'data': 'synthetic',
            'config': {'workload': 'NT-Xent SimCLR loss fwd+bwd, global 2N=8192 (B=4096 pairs, 2 views), D=128, '
                                   'T=0.5, base_T=0.07; rows sharded over %d GPU(s), %d pairs per GPU' % (world, b_local),
                       'engine': engine, 'l2': 'inputs rotate through a pool of %d embedding/gradient sets '
                                               '(> L2 in total); the 2 MiB working set of one step is L2-resident '
                                               'by construction' % pool,
                       'launch': 'public module API (SupConLoss.forward + autograd backward); ' + how},
            'e2e': {'value': B_GLOBAL / (ms_e2e / args.steps * 1e-3), 'unit': 'pairs/s',
                    'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                    'note': 'each step: H2D of its embeddings from pinned memory (prefetched one step ahead on a '
                            'copy stream), forward + backward through the module in graph mode (SupConLoss / '
                            'DistributedSupConLoss with graph=True: one captured forward+backward, exchanges '
                            'included, replayed per step), D2H of the loss, host sync; the gradient stays on the device; '
                            'torch.autograd.set_multithreading_enabled(False) (one device per process)'},
            'gpu_launches': int(launches),
        })
        # roofline of the STEP (6*A*C*D algorithmic FLOPs of SURVEY 8(d) over the graph-replayed step time, per GPU);
        # the isolated kernels are listed under kernel_us
        ach_step = flops_step / world / (per * 1e-3) / 1e12
        out['roofline'] = {'bound': 'tensor', 'achieved': ach_step, 'peak': pk['tf'], 'unit': 'TFLOP/s',
                           'frac': ach_step / pk['tf'],
                           'frac_of_sustained_peak': ach_step / pk['tf_sus'] if pk.get('tf_sus') else None,
                           'traffic': profiled_traffic('supcon_tc_bwd_kernel') if engine != 'fp32' else None,
                           'kernel': 'whole step = supcon_tc_fwd_kernel + supcon_tc_bwd_kernel (%d launches), '
                                     'CUDA-graph replay' % launches,
                           'peak_source': pk['src'] + ' burst',
                           'algorithmic_flops_per_step_per_gpu': flops_step / world}
        if rank == 0 and world == 1:
            kb = loss_kernel_breakdown(engine, min(args.steps, 100))
            t_bwd = kb['backward main kernel']
            out['roofline']['kernel_us'] = kb
            out['roofline']['kernel_us_note'] = ('event-bracketed isolated launches (they include launch gaps and add '
                                                 'up to more than the graph-replayed step)')
            out['roofline']['backward_kernel_frac'] = 4.0 * A * C * DIM / (t_bwd * 1e-6) / 1e12 / pk['tf']
        par = loss_parity(world, rank, engine)
        if rank == 0:
            out['parity'] = par
        if world == 1 and engine == 'bf16':
            # SURVEY 8(d) Metric 1 as the reference calls it: RAW fp32 embeddings, the fused F.normalize, engine
            # 'auto' = the split-bf16 tensor-core engine at the reference's precision (1e-4)
            out['ntxent_fp32'] = loss_fp32_variant(args, pk)
        if rank == 0 and world == 1:
            out['projection_tail'] = head_tail_bench()
        if rank == 0 and not args.no_gpu_baseline:
            out['gpu_baseline'] = gpu_loss_baseline()
            out['gpu_baseline']['ours_over_gpu_baseline_fp32'] = out['value'] / out['gpu_baseline']['fp32']['value']
        barrier(world)
        if world > 1:
            # weak-scaling series of SURVEY 8(d): 512 pairs per GPU, 2N = 1024 * world
            w_res, _, _, _, _ = make_loss_runner(world, rank, engine, 16, b_local=512)
            ms_w, _ = time_steps(w_res, args.steps, args.warmup, world, graph_pool=16)
            out['scaling_weak'] = {'pairs_per_gpu': 512, 'global_2N': 1024 * world, 'ms_per_step': ms_w / args.steps,
                                   'value': 512 * world / (ms_w / args.steps * 1e-3), 'unit': 'pairs/s'}
    # ------------------------------------------------ SyncBN
    if args.workload in ('both', 'syncbn'):
        pool = 12                                    # 12 x (x, dy, y, dx) of 8 MiB each = 400 MB > L2
        step_res, step_e2e, S, bn = make_bn_runner(world, rank, pool)
        c0 = lib.scb_launch_count()
        step_res(0)
        c0 = lib.scb_launch_count()
        step_res(1)
        launches = lib.scb_launch_count() - c0
        ms, how = time_steps(step_res, args.steps, args.warmup, world, graph_pool=pool)
        # one device per process: autograd may run the backward in the calling thread instead of handing it to its
        # device thread (a PyTorch switch any training script can set; 55 -> 39 us of host time per backward)
        with torch.autograd.set_multithreading_enabled(False):
            ms_e2e, _ = time_steps(step_e2e, args.steps, args.warmup, world)
        per = ms / args.steps
        gbs = 8.0 * S / (per * 1e-3) / 1e9           # per GPU; algorithmic bytes of SURVEY 8(d)
        sb = {'metric': 'SyncBN fwd+bwd HBM GB/s', 'value': gbs * world, 'per_gpu': gbs, 'unit': 'GB/s',
              'ms_per_step': per, 'scaling': 'weak', 'dtype': 'bf16 I/O, f32 statistics',
              'config': {'workload': 'SyncBatchNorm fwd+bwd on [256,256,8,8] NHWC bf16 per GPU, %d GPU(s)' % world,
                         'l2': 'x/dy/y/dx rotate through %d buffer sets (> L2)' % pool,
                         'launch': 'public module API (SyncBatchNorm.forward + autograd backward); ' + how},
              'e2e': {'value': 8.0 * S * world / (ms_e2e / args.steps * 1e-3) / 1e9, 'unit': 'GB/s',
                      'h2d_bytes_per_step': 2 * S, 'd2h_bytes_per_step': 2 * S,
                      'note': 'bound by the host link: x and dy go in, y and dx come back, 4*S = 33.6 MB per step '
                              'over PCIe against 23 us of kernels -- a PCIe figure, not a kernel one'},
              'gpu_launches': int(launches)}
        # roofline of the STEP: 8*S algorithmic bytes over the graph-replayed fwd+bwd, per GPU
        sb['roofline'] = {'bound': 'hbm', 'achieved': gbs, 'peak': pk['hbm'], 'unit': 'GB/s', 'frac': gbs / pk['hbm'],
                          'traffic': (profiled_traffic('bn_bwd_cluster_smem') or profiled_traffic('bn_bwd_cluster')
                                      or profiled_traffic('bn_fwd_cluster')),
                          'kernel': 'whole step (%d launches): '
                          'bn_fwd_cluster_kernel + bn_bwd_cluster_smem_kernel; traffic = the backward kernel' % launches,
                          'peak_source': pk['src'], 'algorithmic_bytes_per_step': 8 * S,
                          'note': 'the single-launch passes move 5*S (x once forward; x, dy once backward; y, dx written)'}
        if rank == 0 and world == 1:
            kb = bn_kernel_breakdown(min(args.steps, 100))
            sb['roofline']['kernel_us'] = kb
            sb['roofline']['kernel_us_note'] = 'event-bracketed isolated launches (include launch gaps)'
        parb = bn_parity(world, rank)
        if rank == 0:
            sb['parity'] = parb
        if not args.no_gpu_baseline:
            gb = gpu_bn_baseline(world)
            if rank == 0:
                sb['gpu_baseline'] = gb
                sb['gpu_baseline']['ours_over_gpu_baseline'] = gbs / gb['value']
        if 'metric' in out:
            out['syncbn'] = sb
        else:
            out.update(sb)
            out.update({'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup, 'higher_is_better': True,
                        'vs_baseline': None, 'data': 'synthetic'})

    # keep the GPU under the same load until the clock sampler has a few readings.  Every rank runs
    # the SAME number of extra steps (they contain collectives when N > 1): derived from the
    # max-over-ranks step time, which all ranks hold.
    n_extra = max(20, min(20000, int(1200.0 / max(ms / args.steps, 1e-3))))
    for i in range(n_extra):
        step_res(i)
        if i % 50 == 49:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    if rank == 0:
        out['clocks'] = sampler.stop(t_load0, time.time())
        out['clocks']['note'] = ('sampled every 100 ms from the first warm-up step to ~1.2 s of the same steps '
                                 'run right after the timed regions (the timed regions alone are shorter than '
                                 'one sampling period)')
    barrier(world)

    # ------------------------------------------------ CPU baseline (rank 0, N=1)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        t, n, cores = cpu_loss_baseline()
        out['cpu_baseline'] = {'value': B_GLOBAL / t, 'unit': 'pairs/s', 'cores': cores, 'kind': 'port',
                               'sample': 'median of %d fwd+bwd of the full [4096,2,128] fp32 workload '
                                         '(oracle/dense_port.py, torch CPU)' % n}
        if 'syncbn' in out:
            tb, nb, _ = cpu_bn_baseline()
            S = BN_SHAPE[0] * BN_SHAPE[1] * BN_SHAPE[2] * BN_SHAPE[3] * 2
            out['syncbn']['cpu_baseline'] = {'value': 8 * S / tb / 1e9, 'unit': 'GB/s', 'cores': cores,
                                             'kind': 'port', 'sample': 'median of %d fwd+bwd, F.batch_norm '
                                                                       'train mode, NHWC bf16, torch CPU' % nb}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
