"""``python -m dmlp_b200.grpc_client`` — the reference client's training loop, spoken entirely over the gRPC contract.

The reference client (``DSML/client/client.go:516-659``) computes forward / backward itself on the CPU, copies the SAME
gradient to every device, calls ``AllReduceRing``, reads device 1 back, applies SGD and re-uploads the weights. Here
the client only moves data and issues RPCs; every number is produced where the buffers live:

    per step, for every device d (its own shard of the batch stream — real data parallelism):
        Memcpy H2D   data  -> 0x3000, labels -> 0x4000              (client.go:29-32 memory map)
        RunForward   0x3000 -> logits at 0x5000                      (loss / accuracy for the log line)
        RunBackward  gradients -> 0x1000
    AllReduceRing(count = blob bytes) with metadata dmlp-sgd-lr / dmlp-sgd-weights=0x2000:
        ring / one-shot kernel over peer memory that averages the gradients and applies SGD to the weights at
        0x2000 on every device in the same kernel (``backend=cuda``); the portable ring + update on ``sim``

Works against ``serve_coordinator --with-devices`` or any coordinator whose devices were started separately.
"""
from __future__ import annotations

import argparse
import logging
import time
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional, Sequence

import numpy as np

from .models.layout import ParamLayout
from .proto import gpu_sim as pb
from .proto import rpc
from .utils import data as D

log = logging.getLogger("dmlp.grpc_client")
GRAD_ADDR, WEIGHT_ADDR, DATA_ADDR, LABEL_ADDR, LOGIT_ADDR = 0x1000, 0x2000, 0x3000, 0x4000, 0x5000


class GrpcTrainer:
    """Data-parallel SGD driven through GPUCoordinator (communicator, Memcpy, AllReduceRing) and GPUDevice
    (RunForward / RunBackward)."""

    def __init__(self, coordinator_addr: str, device_addrs: Sequence[str], dims: Sequence[int] = (784, 128, 64, 10),
                 batch: int = 64, lr: float = 0.01, seed: int = 0):
        self.dims, self.batch, self.lr = tuple(dims), batch, lr
        self.layout = ParamLayout(dims)
        self.coord = rpc.CoordinatorStub(rpc.insecure_channel(coordinator_addr))
        self.dev_stubs = [rpc.DeviceStub(rpc.insecure_channel(a)) for a in device_addrs]
        init = self.coord.CommInit(pb.CommInitRequest(numDevices=len(device_addrs), device_addresses=list(device_addrs)))
        if not init.success:
            raise RuntimeError("CommInit failed")
        self.comm_id = init.commId
        self.device_ids = [m.deviceId.value for m in init.devices]
        self._stub = dict(zip(self.device_ids, self.dev_stubs))
        # the per-device halves of a step are independent RPC sequences: issue them concurrently (gRPC releases the GIL
        # while it waits; the reference client walks its devices one after the other, client.go:596-647)
        self._pool = ThreadPoolExecutor(max_workers=max(len(self.device_ids), 1), thread_name_prefix="dmlp-dev")
        self.blob_bytes = self.layout.reference_count * 4
        blob = self.layout.to_reference_bytes(self.layout.init_reference(seed))
        for did in self.device_ids:  # identical replicas (client.go:642-644)
            self._h2d(did, WEIGHT_ADDR, blob)

    # ------------------------------------------------------------------ plumbing
    # (GPUDevice.Memcpy, not the coordinator's: the coordinator keeps the reference's hard-coded 0x1000..0x2000 window,
    #  gpu_coordinator_server.go:199-201, while the device checks against its real memory size)
    def _h2d(self, did: int, addr: int, data: bytes) -> None:
        self._stub[did].Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
            hostSrcData=data, dstDeviceId=pb.DeviceId(value=did), dstMemAddr=pb.MemAddr(value=addr))))

    def _d2h(self, did: int, addr: int) -> bytes:
        return self._stub[did].Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
            srcDeviceId=pb.DeviceId(value=did), srcMemAddr=pb.MemAddr(value=addr), numBytes=0))).deviceToHost.dstData

    def weights_blob(self, device_index: int = 0) -> bytes:
        return self._d2h(self.device_ids[device_index], WEIGHT_ADDR)

    # ------------------------------------------------------------------ one optimisation step
    def _device_half(self, did: int, stub, x, y, backward: bool = True):
        """Everything one device does for one batch: upload its shard, RunForward (+ RunBackward), read the logits."""
        classes = self.dims[-1]
        self._h2d(did, DATA_ADDR, np.ascontiguousarray(x, dtype="<f4").tobytes())
        if backward:
            self._h2d(did, LABEL_ADDR, np.eye(classes, dtype="<f4")[np.asarray(y)].tobytes())  # one-hot like client.go:315-350
        stub.RunForward(pb.RunForwardRequest(deviceId=did, inputAddr=DATA_ADDR, outputAddr=LOGIT_ADDR))
        if backward:
            stub.RunBackward(pb.RunBackwardRequest(deviceId=did, gradientAddr=GRAD_ADDR))
        logits = np.frombuffer(self._d2h(did, LOGIT_ADDR), dtype="<f4").reshape(len(y), classes).astype(np.float64)
        z = logits - logits.max(1, keepdims=True)
        p = np.exp(z) / np.exp(z).sum(1, keepdims=True)
        loss = float(-np.log(p[np.arange(len(y)), y] + 1e-10).sum())
        return loss, int((logits.argmax(1) == np.asarray(y)).sum()), len(y)

    def step(self, shards) -> dict:
        """``shards[d] = (x float32 [B][in], y int [B])`` for device d."""
        futs = [self._pool.submit(self._device_half, did, stub, x, y)
                for did, stub, (x, y) in zip(self.device_ids, self.dev_stubs, shards)]
        parts = [f.result() for f in futs]
        loss_sum, correct, seen = (sum(p[i] for p in parts) for i in range(3))
        r = self.coord.AllReduceRing(pb.AllReduceRingRequest(commId=self.comm_id, count=self.blob_bytes),
                                     metadata=(("dmlp-dtype", "f32"), ("dmlp-sgd-lr", repr(self.lr)),
                                               ("dmlp-sgd-weights", hex(WEIGHT_ADDR))))
        if not r.success:
            raise RuntimeError("AllReduceRing failed")
        return {"loss": loss_sum / seen, "accuracy": correct / seen}

    def evaluate(self, x: np.ndarray, y: np.ndarray) -> dict:
        """Forward-only pass over all full batches, spread over the devices (the reference's ``testModel``,
        client.go:462-502: softmax + argmax accuracy, "Final Test Accuracy")."""
        B, n = self.batch, len(self.device_ids)
        nb = x.shape[0] // B
        loss_sum, correct, seen = 0.0, 0, 0
        for b0 in range(0, nb, n):
            futs = [self._pool.submit(self._device_half, self.device_ids[d], self.dev_stubs[d],
                                      x[(b0 + d) * B:(b0 + d + 1) * B], y[(b0 + d) * B:(b0 + d + 1) * B], False)
                    for d in range(min(n, nb - b0))]
            for f in futs:
                ls, c, m = f.result()
                loss_sum, correct, seen = loss_sum + ls, correct + c, seen + m
        acc = correct / max(seen, 1)
        log.info("Final Test Accuracy: %.2f%%", 100 * acc)  # client.go:501
        return {"loss": loss_sum / max(seen, 1), "accuracy": acc, "samples": seen}

    def fit(self, x: np.ndarray, y: np.ndarray, steps: int, first_batch: int = 0) -> List[dict]:
        n, B = len(self.device_ids), self.batch
        nb = x.shape[0] // B
        out = []
        for s in range(steps):
            shards = []
            for d in range(n):  # device d takes batch s*n + d of the stream (parallel.trainer.shard_schedule)
                b = (first_batch + s * n + d) % nb
                shards.append((x[b * B:(b + 1) * B], y[b * B:(b + 1) * B]))
            out.append(self.step(shards))
        return out

    def close(self) -> None:
        self._pool.shutdown(wait=True)
        self.coord.CommDestroy(pb.CommDestroyRequest(commId=self.comm_id))


def main(argv: Optional[Sequence[str]] = None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    # the reference client's endpoints (client.go:518,527) = this repo's launchers' defaults (services/server.py)
    ap.add_argument("--coordinator", default="127.0.0.1:50051")
    ap.add_argument("--devices", default="127.0.0.1:5003,127.0.0.1:5004,127.0.0.1:5005")
    ap.add_argument("--dims", default="784,128,64,10")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--steps-per-epoch", type=int, default=50)
    ap.add_argument("--samples", type=int, default=60000)
    ap.add_argument("--test-samples", type=int, default=10000)
    a = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(message)s")
    dims = tuple(int(v) for v in a.dims.split(","))
    t = GrpcTrainer(a.coordinator, a.devices.split(","), dims=dims, batch=a.batch, lr=a.lr)
    x, y = D.synthetic_mnist(a.samples, seed=0, features=dims[0], num_classes=dims[-1])
    log.info("Loaded %d training samples", x.shape[0])
    log.info("Starting MLP training...")
    step0 = 0
    for ep in range(1, a.epochs + 1):
        t0 = time.time()
        st = t.fit(x, y, a.steps_per_epoch, first_batch=step0)
        step0 += a.steps_per_epoch * len(t.device_ids)
        log.info("Epoch %d complete: Avg Loss: %.4f, Accuracy: %.2f%%  (%.1f steps/s)", ep,
                 float(np.mean([s["loss"] for s in st])), 100 * float(np.mean([s["accuracy"] for s in st])),
                 a.steps_per_epoch / (time.time() - t0))
    log.info("Training complete.")
    xte, yte = D.synthetic_mnist(a.test_samples, seed=1, features=dims[0], num_classes=dims[-1])
    t.evaluate(xte, yte)  # testModel (client.go:658 -> :462-502)
    t.close()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
