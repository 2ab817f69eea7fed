"""Multi-process test harness: run ``fn(rank, world, *args)`` in ``world`` fresh processes with a
torchrun-like environment (gloo on CPU).  The reference has no such tier - everything distributed there
needs MPI and GPUs; this is what lets pipeline / tensor / data parallel plumbing run in CI."""

import os
import socket
import sys
import traceback

import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _to_wire(obj):
    """Tensors cross the process boundary as numpy arrays (torch's fd-passing breaks when the child exits)."""
    import torch

    if torch.is_tensor(obj):
        return ("__tensor__", obj.detach().cpu().float().numpy(), str(obj.dtype))
    if isinstance(obj, dict):
        return {k: _to_wire(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_wire(v) for v in obj)
    return obj


def _from_wire(obj):
    import torch

    if isinstance(obj, tuple) and len(obj) == 3 and obj[0] == "__tensor__":
        return torch.from_numpy(obj[1])
    if isinstance(obj, dict):
        return {k: _from_wire(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_from_wire(v) for v in obj)
    return obj


def _orderly_shutdown(success):
    mod = sys.modules.get("smdistributed_modelparallel_b200")
    try:
        if mod is not None and mod.state.core is not None:
            mod.state.core.shutdown(success)
    except Exception:                        # noqa: BLE001 - teardown must not mask the test's own result
        pass


def _entry(rank, world, port, fn, args, queue, use_cuda=False, env=None):
    if env:                                  # a fork-server child starts from the server's environment
        os.environ.update(env)
    os.environ.update({"RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank),
                       "LOCAL_WORLD_SIZE": str(world), "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
                       "SMP_LOG_LEVEL": os.environ.get("SMP_LOG_LEVEL", "warning")})
    if not use_cuda:
        os.environ["CUDA_VISIBLE_DEVICES"] = ""
        import torch
        # tiny tensors, several ranks per machine: intra-op threads only fight one another
        torch.set_num_threads(max(1, (os.cpu_count() or 4) // (2 * world)))
    sys.path.insert(0, ROOT)
    import faulthandler
    faulthandler.enable()                    # a crash inside native code leaves the Python stack of every thread
    dump_after = float(os.environ.get("SMP_TEST_HANG_DUMP_S", "0"))
    if dump_after > 0:                       # where is every thread if the worker is still running by then?
        import faulthandler
        faulthandler.dump_traceback_later(dump_after, exit=os.environ.get("SMP_TEST_HANG_DUMP_NOEXIT", "0") != "1")
    try:
        out = _to_wire(fn(rank, world, *args))
        queue.put((rank, "ok", out))
    except Exception:
        queue.put((rank, "err", traceback.format_exc()))
        raise
    finally:
        # a multiprocessing child leaves through os._exit: no atexit handlers, so the library's orderly shutdown
        # (flush pending control messages, tell the peers goodbye) has to be called here - without it the peers see
        # this rank's sockets close unannounced, which the message bus reports as a lost peer
        _orderly_shutdown(sys.exc_info()[0] is None)


def run_distributed(fn, world, *args, timeout=240, use_cuda=False):
    timeout = float(os.environ.get("SMP_TEST_TIMEOUT_S", timeout))
    # CPU workers are forked from a server that has already imported torch (seconds per process otherwise);
    # the server never touches torch.distributed or CUDA, so its children start clean.  GPU workers are spawned.
    if use_cuda or os.environ.get("SMP_TEST_START_METHOD") == "spawn":
        ctx = mp.get_context("spawn")
    else:
        ctx = mp.get_context("forkserver")
        # the HuggingFace modeling modules cost seconds to import; the server pays that once for every worker of the
        # session (a module that is not installed is skipped by the fork server)
        ctx.set_forkserver_preload(["torch", "numpy", "transformers", "transformers.models.gpt2.modeling_gpt2",
                                    "transformers.models.gptj.modeling_gptj",
                                    "transformers.models.gpt_neox.modeling_gpt_neox",
                                    "transformers.models.gpt_neo.modeling_gpt_neo",
                                    "transformers.models.bert.modeling_bert"])
    q = ctx.Queue()
    port = _free_port()
    env = dict(os.environ)
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, q, use_cuda, env)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    try:
        import queue
        import time
        deadline = time.monotonic() + timeout
        while len(results) < world:
            try:
                rank, status, payload = q.get(timeout=1.0)
            except queue.Empty:
                # a worker that died without reporting (a crash in native code, a kill) will never answer: say which
                # one and how it ended instead of waiting out the whole timeout
                dead = [(r, p.exitcode) for r, p in enumerate(procs)
                        if p.exitcode not in (None, 0) and r not in results]
                if dead:
                    time.sleep(1.0)                      # let a result that was already on its way arrive
                    if q.empty():
                        raise AssertionError(f"worker(s) died without a result (rank, exit code): {dead}") from None
                if time.monotonic() > deadline:
                    raise AssertionError(f"no result from ranks {sorted(set(range(world)) - set(results))} "
                                         f"within {timeout:.0f} s") from None
                continue
            if status != "ok":
                raise AssertionError(f"rank {rank} failed:\n{payload}")
            results[rank] = _from_wire(payload)
    finally:
        for p in procs:
            p.join(timeout=20)
            if p.is_alive():
                p.kill()
    return [results[r] for r in range(world)]
