"""Training entry point: the reference's ``resnet/main.py`` re-hosted on the B200-native stack.

Launch exactly like the reference:

    torchrun --nproc_per_node=8 resnet/main.py [flags]
    python -m torch.distributed.launch --nproc_per_node=8 resnet/main.py [flags]     (passes --local-rank)

Flags, defaults, prints, evaluation / checkpoint cadence and checkpoint format are the reference's
(reference resnet/main.py:40-124); the documented defects of the shipped file are fixed (the
SyntaxError at :111, the ``model_filename`` KeyError at :58, the undefined ``accuracy`` at :114, ``--learning_rate``
typed int at :55, ``--local_rank`` only / no LOCAL_RANK fallback at :52, model_dir never created).  Extra switches
select the other BASELINE.json configurations (model, image size, synthetic data, SyncBN, bucket cap, backend).
"""
import argparse
import os
import time

import torch
import torch.distributed as dist

from . import optim
from .models.resnet import MODELS
from .ops.loss import CrossEntropyLoss
from .parallel import (DistributedDataParallel, DistributedSampler, convert_sync_batchnorm, destroy_process_group,
                       init_process_group)
from .utils.checkpoint import load_checkpoint, save_checkpoint
from .utils.data import DevicePrefetcher, SyntheticImages, cifar10_datasets, make_loaders
from .utils.seed import set_random_seeds


def evaluate(model, device, test_loader):
    """Rank-local accuracy over a loader (reference resnet/main.py:23-37). Issues no collectives."""
    model.eval()
    correct = 0
    total = 0
    with torch.no_grad():
        for data in test_loader:
            images, labels = data[0].to(device), data[1].to(device)
            outputs = model(images)
            _, predicted = torch.max(outputs.data, 1)
            total += labels.size(0)
            correct += (predicted == labels).sum().item()
    return correct / total


def build_parser():
    defaults = {"num_epochs": 10000, "batch_size": 256, "lr": 0.01, "seed": 0, "model_dir": "saved_models",
                "model_filename": "resnet_distributed.pth"}
    p = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    p.add_argument("--local_rank", "--local-rank", dest="local_rank", type=int, default=None,
                   help="Local rank. Passed by torch.distributed.launch; torchrun sets LOCAL_RANK instead")
    p.add_argument("--num_epochs", type=int, help="Number of training epochs", default=defaults["num_epochs"])
    p.add_argument("--batch-size", type=int, help="Training batch size (per rank)", default=defaults["batch_size"])
    p.add_argument("--learning_rate", type=float, help="Learning rate", default=defaults["lr"])
    p.add_argument("--seed", type=int, help="Random seed for training", default=defaults["seed"])
    p.add_argument("--model_dir", type=str, help="Model directory to store saved models",
                   default=defaults["model_dir"])
    p.add_argument("--model_filename", type=str, help="Model filename to be saved",
                   default=defaults["model_filename"])
    p.add_argument("--resume", action="store_true", help="Resume training from saved checkpoint.")
    # ---- beyond the reference
    p.add_argument("--model", default="resnet18", choices=sorted(MODELS))
    p.add_argument("--backend", default=None, help="gloo (CPU plumbing) | b200/nccl (GPU, P2P kernels)")
    p.add_argument("--synthetic", action="store_true", help="synthetic data of the dataset's shape")
    p.add_argument("--image-size", type=int, default=32)
    p.add_argument("--num-classes-data", type=int, default=10)
    p.add_argument("--dataset-size", type=int, default=50000)
    p.add_argument("--data-root", default="data")
    p.add_argument("--num-workers", type=int, default=8)
    p.add_argument("--bucket-cap-mb", type=float, default=None)
    p.add_argument("--sync-bn", action="store_true")
    p.add_argument("--no-fuse-optimizer", action="store_true",
                   help="run SGD in optimizer.step() instead of as the all-reduce epilogue")
    p.add_argument("--eval-every", type=int, default=10)
    p.add_argument("--max-steps", type=int, default=0, help="stop after this many optimizer steps (0 = no limit)")
    p.add_argument("--reshuffle-each-epoch", action="store_true", help="call sampler.set_epoch (the reference never does)")
    p.add_argument("--log-every", type=int, default=0, help="print loss / images-per-second every N steps")
    p.add_argument("--cuda-graph", action="store_true",
                   help="capture the whole training step into a CUDA graph (one launch per step)")
    return p


def main(argv=None):
    args = build_parser().parse_args(argv)
    local_rank = args.local_rank if args.local_rank is not None else int(os.environ.get("LOCAL_RANK", "0"))
    model_filepath = os.path.join(args.model_dir, args.model_filename)
    set_random_seeds(args.seed)

    init_process_group(backend=args.backend, local_rank=local_rank)
    use_cuda = torch.cuda.is_available() and args.backend != "gloo"
    device = torch.device("cuda:{}".format(local_rank)) if use_cuda else torch.device("cpu")

    model = MODELS[args.model](pretrained=False)
    if args.sync_bn:
        model = convert_sync_batchnorm(model)
    model = model.to(device)
    ddp_model = DistributedDataParallel(model, device_ids=[local_rank] if use_cuda else None,
                                        output_device=local_rank if use_cuda else None,
                                        bucket_cap_mb=args.bucket_cap_mb)
    if args.resume:
        load_checkpoint(ddp_model, model_filepath, local_rank=local_rank if use_cuda else None)

    if args.synthetic:
        shape = (3, args.image_size, args.image_size)
        train_set = SyntheticImages(args.dataset_size, shape, args.num_classes_data, seed=args.seed)
        test_set = SyntheticImages(max(args.dataset_size // 5, 1), shape, args.num_classes_data, seed=args.seed + 1)
    else:
        train_set, test_set = cifar10_datasets(args.data_root)
    train_sampler = DistributedSampler(dataset=train_set)
    train_loader, test_loader = make_loaders(train_set, test_set, args.batch_size, train_sampler,
                                             num_workers=args.num_workers, pin_memory=use_cuda)

    criterion = CrossEntropyLoss()
    optimizer = optim.SGD(ddp_model.parameters(), lr=args.learning_rate, momentum=0.9, weight_decay=1e-5,
                          fuse_with_backward=use_cuda and not args.no_fuse_optimizer)

    graphed = None
    if args.cuda_graph and use_cuda:
        from .utils.graph import GraphedTrainStep
        graphed = GraphedTrainStep(ddp_model, criterion, optimizer)

    steps = 0
    for epoch in range(args.num_epochs):
        print("Local Rank: {}, Epoch: {}, Training ...".format(local_rank, epoch), flush=True)
        if epoch % args.eval_every == 0:
            if local_rank == 0:
                acc = evaluate(model=ddp_model, device=device, test_loader=test_loader)
                save_checkpoint(ddp_model, model_filepath)
                print("-" * 75)
                print("Epoch: {}, Accuracy: {}".format(epoch, acc))
                print("-" * 75, flush=True)
            if dist.is_initialized() and dist.get_world_size() > 1:
                # The other ranks wait here, on the host, for rank 0's evaluation + checkpoint (the reference makes
                # them wait inside their first NCCL collective instead, resnet/main.py:109-115). A CPU-tensor
                # all-reduce always travels over the gloo control plane; it is bounded by the process-group timeout
                # and keeps the ranks out of the device-side spin loops while nothing can answer them.
                dist.all_reduce(torch.zeros(1))
        ddp_model.train()
        if args.reshuffle_each_epoch:
            train_sampler.set_epoch(epoch)
        prefetch = use_cuda and os.environ.get("B200DDP_NO_PREFETCH", "0") != "1"
        batches = DevicePrefetcher(train_loader, device) if prefetch else train_loader
        t0, seen = time.perf_counter(), 0
        for data in batches:
            inputs, labels = data[0].to(device), data[1].to(device)
            if graphed is not None:
                loss = graphed(inputs, labels)
            else:
                optimizer.zero_grad()
                outputs = ddp_model(inputs)
                loss = criterion(outputs, labels)
                loss.backward()
                optimizer.step()
            steps += 1
            seen += inputs.shape[0]
            if args.log_every and steps % args.log_every == 0:
                lv = loss.item()
                dt = time.perf_counter() - t0
                print("rank {} step {} loss {:.4f} {:.1f} img/s/rank".format(local_rank, steps, lv, seen / dt),
                      flush=True)
                t0, seen = time.perf_counter(), 0
            if args.max_steps and steps >= args.max_steps:
                break
        if args.max_steps and steps >= args.max_steps:
            break
    if use_cuda:
        torch.cuda.synchronize()
    destroy_process_group()
    return 0


if __name__ == "__main__":
    main()
