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
    def step(self, shards) -> dict:
        """``shards[d] = (x float32 [B][in], y int [B])`` for device d."""
        classes = self.dims[-1]
        loss_sum, correct, seen = 0.0, 0, 0
        for did, stub, (x, y) in zip(self.device_ids, self.dev_stubs, shards):
            self._h2d(did, DATA_ADDR, np.ascontiguousarray(x, dtype="<f4").tobytes())
            self._h2d(did, LABEL_ADDR, np.eye(classes, dtype="<f4")[np.asarray(y)].tobytes())  # one-hot like client.go:315-350
            stub.RunForward(pb.RunForwardRequest(deviceId=did, inputAddr=DATA_ADDR, outputAddr=LOGIT_ADDR))
            stub.RunBackward(pb.RunBackwardRequest(deviceId=did, gradientAddr=GRAD_ADDR))
            logits = np.frombuffer(self._d2h(did, LOGIT_ADDR), dtype="<f4").reshape(len(y), classes).astype(np.float64)
            z = logits - logits.max(1, keepdims=True)
            p = np.exp(z) / np.exp(z).sum(1, keepdims=True)
            loss_sum += float(-np.log(p[np.arange(len(y)), y] + 1e-10).sum())
            correct += int((logits.argmax(1) == np.asarray(y)).sum())
            seen += len(y)
        r = self.coord.AllReduceRing(pb.AllReduceRingRequest(commId=self.comm_id, count=self.blob_bytes),
                                     metadata=(("dmlp-dtype", "f32"), ("dmlp-sgd-lr", repr(self.lr)),
                                               ("dmlp-sgd-weights", hex(WEIGHT_ADDR))))
        if not r.success:
            raise RuntimeError("AllReduceRing failed")
        return {"loss": loss_sum / seen, "accuracy": correct / seen}

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
        self.coord.CommDestroy(pb.CommDestroyRequest(commId=self.comm_id))


def main(argv: Optional[Sequence[str]] = None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--coordinator", default="127.0.0.1:50050")
    ap.add_argument("--devices", default="127.0.0.1:50051,127.0.0.1:50052,127.0.0.1:50053")
    ap.add_argument("--dims", default="784,128,64,10")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--steps-per-epoch", type=int, default=50)
    ap.add_argument("--samples", type=int, default=60000)
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
    t.close()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
