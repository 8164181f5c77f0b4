"""``python -m dmlp_b200.train`` — the reference client's ``main()`` on real GPUs.

Reference flow (``DSML/client/client.go:516-659``): connect, CommInit, load MNIST, 10 epochs of
batch-64 SGD with an all-reduce per batch, per-epoch "Avg Loss / Accuracy" lines, final test accuracy.
Same flow here, one process per GPU (launch with ``torch.distributed.run`` for more than one), data
parallel with the all-reduce fused into the SGD update. Real MNIST is used when the IDX files exist
(``--synthetic false``); the reference ships only the t10k split, so the default is synthetic
MNIST-shaped data for training and, when present, the real t10k images for a sanity evaluation.
"""
from __future__ import annotations

import argparse
import logging
import os
import sys
import time

from .utils import data as D
from .utils.config import Config
from .utils.metrics import MetricsWriter, setup_logging

log = logging.getLogger("dmlp.train")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    Config.add_arguments(ap)
    ap.add_argument("--steps-per-epoch", type=int, default=0)
    ap.add_argument("--checkpoint", default="", help="write <path> (reference weight blob) + <path>.json after every epoch")
    ap.add_argument("--resume", default="", help="checkpoint to continue from (weights, step counter, epoch)")
    cfg = Config.from_args(ap.parse_args(argv))
    ns = ap.parse_args(argv)
    setup_logging()
    import torch
    import torch.distributed as dist
    from .parallel import schedule as S
    from .parallel.trainer import Trainer, TrainerConfig, pin
    from .runtime import native

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    found = D.find_reference_data(cfg.data_dir) or {}
    if not cfg.synthetic and "train_images" in found and "train_labels" in found:
        xtr, ytr = D.load_idx_images(found["train_images"]), D.load_idx_labels(found["train_labels"])
    else:
        if not cfg.synthetic and rank == 0:
            log.info("train images not found under %s; using synthetic MNIST-shaped data", cfg.data_dir)
        xtr, ytr = D.synthetic_mnist(cfg.train_samples, seed=cfg.seed, features=cfg.dims[0], num_classes=cfg.dims[-1])
    if rank == 0:
        log.info("Loaded %d training samples", xtr.shape[0])
    xte, yte = D.synthetic_mnist(cfg.test_samples, seed=cfg.seed + 1, features=cfg.dims[0], num_classes=cfg.dims[-1])
    if rank == 0:
        log.info("Loaded %d test samples", xte.shape[0])

    if cfg.allreduce_table:
        from .runtime.comm import set_allreduce_table
        set_allreduce_table(cfg.allreduce_table)
    tcfg = TrainerConfig(dims=cfg.dims, batch=cfg.batch, lr=cfg.lr, precision=cfg.precision, allreduce=cfg.allreduce,
                         overlap=cfg.overlap, seed=cfg.seed, watchdog_ms=cfg.watchdog_ms)
    trainer = Trainer(tcfg, device=local, process_group=world > 1)
    if cfg.weights_in:
        trainer.load_weights(cfg.weights_in)
    first_epoch = 1
    if ns.resume:
        meta = trainer.load_checkpoint(ns.resume)  # every rank reads the same file: replicas start identical
        first_epoch = int(meta.get("epoch", 0)) + 1
        if rank == 0:
            log.info("resumed from %s: epoch %d, step %d", ns.resume, first_epoch - 1, trainer.global_step)
    # fault injection (BASELINE config 5): "dead-rank:<r>@step:<s>" - rank r stops taking part in the gradient exchange
    # at global step s, as if its GPU had died (reference test: gpu_coordinator_server_test.go:370-429)
    dead_rank, dead_step = S.parse_fault(cfg.fault)
    if dead_rank == rank:
        trainer.inject_fault(dead_step)
    px, py = pin(xtr), pin(ytr)
    spe = ns.steps_per_epoch or xtr.shape[0] // (cfg.batch * world)
    sched = S.make(cfg.schedule, cfg.lr, total_steps=spe * cfg.epochs)
    mw = MetricsWriter(cfg.metrics_jsonl if rank == 0 else None)
    if rank == 0:
        log.info("Starting MLP training...")
    for ep in range(first_epoch, cfg.epochs + 1):
        try:
            st = trainer.fit(px, py, spe, lr_schedule=sched, device_timing=True)
        except native.NativeError as e:
            if e.code not in (native.ERR_TIMEOUT, native.ERR_ABORTED):
                raise
            # A peer stopped answering mid-all-reduce: the in-kernel watchdog fired, the communicator is FAILED (sticky,
            # like the reference's, gpu_coordinator_server.go:339-346). Log it, leave no GPU work behind, exit non-zero.
            log.error("rank %d: communicator FAILED in epoch %d: %s", rank, ep, e.message)
            mw.write("failure", rank=rank, epoch=ep, error=e.message, global_step=trainer.global_step)
            mw.close()
            trainer.close()
            return 3
        if dead_rank == rank and trainer.global_step > dead_step >= 0:
            # the "dead" rank: its kernels went silent at the fault step. Give the survivors their watchdog budget to
            # notice before this process disappears (a launcher that sees a child exit tears the others down).
            log.error("rank %d: simulated device failure at step %d, exiting", rank, dead_step)
            time.sleep(cfg.watchdog_ms / 1000.0 + 1.0)
            mw.close()
            return 17
        if rank == 0:
            log.info("Epoch %d complete: Avg Loss: %.4f, Accuracy: %.2f%%", ep, st["loss"], 100 * st["accuracy"])
            mw.write("epoch", epoch=ep, loss=st["loss"], accuracy=st["accuracy"], steps=spe,
                     device_ms_per_step=st["device_ms"] / spe, samples_per_s=spe * cfg.batch * world / (st["device_ms"] / 1e3))
            if ns.checkpoint:
                trainer.save_checkpoint(ns.checkpoint, epoch=ep, schedule=cfg.schedule)
    if rank == 0:
        log.info("Training complete.")
    ev = trainer.evaluate(xte, yte)
    if rank == 0:
        mw.write("eval", **ev)
    if "test_images" in found and "test_labels" in found and cfg.dims[0] == 784 and cfg.dims[-1] == 10:
        # the reference ships the real t10k split: evaluate on it too (meaningful when the run trained on real data)
        xr, yr = D.load_idx_images(found["test_images"]), D.load_idx_labels(found["test_labels"])
        evr = trainer.evaluate(xr, yr, label="Accuracy on the reference's t10k images")
        if rank == 0:
            mw.write("eval_t10k", **evr)
    if rank == 0 and cfg.weights_out:
        trainer.save_weights(cfg.weights_out)
        log.info("weights saved to %s (reference client layout, %d bytes)", cfg.weights_out,
                 os.path.getsize(cfg.weights_out))
    mw.close()
    trainer.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
