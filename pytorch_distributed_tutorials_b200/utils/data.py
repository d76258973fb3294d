"""Data pipeline pieces: the reference's CIFAR-10 transforms/loaders (reference resnet/main.py:87-100), a synthetic
dataset of the same shape for machines without the data, and a pinned-memory device prefetcher that overlaps the
host->device copy of the next batch with the current step (the reference does a blocking pageable ``.to(device)``
per batch, resnet/main.py:119)."""
import torch
from torch.utils.data import DataLoader, Dataset

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)


class SyntheticImages(Dataset):
    """Deterministic random images/labels with a given shape (CIFAR-10: 3x32x32, 10 classes)."""

    def __init__(self, n=50000, shape=(3, 32, 32), num_classes=10, seed=0):
        g = torch.Generator().manual_seed(seed)
        self.n, self.shape, self.num_classes = n, shape, num_classes
        self.labels = torch.randint(0, num_classes, (n,), generator=g)
        self.seed = seed

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 1000003 + i)
        return torch.randn(self.shape, generator=g), int(self.labels[i])


def cifar10_datasets(root="data"):
    """Train and test CIFAR-10 with the reference's transform (the *augmenting* transform is used for the test
    set too, as in the reference, resnet/main.py:87-95). ``download=False``: the data must already exist."""
    import torchvision
    import torchvision.transforms as T
    tf = T.Compose([T.RandomCrop(32, padding=4), T.RandomHorizontalFlip(), T.ToTensor(),
                    T.Normalize(CIFAR_MEAN, CIFAR_STD)])
    train = torchvision.datasets.CIFAR10(root=root, train=True, download=False, transform=tf)
    test = torchvision.datasets.CIFAR10(root=root, train=False, download=False, transform=tf)
    return train, test


def make_loaders(train_set, test_set, batch_size, sampler, num_workers=8, pin_memory=False):
    """Loader arguments of the reference (resnet/main.py:98,100): per-rank batch, 8 workers, test batch 128,
    no drop_last (ragged last batch)."""
    train = DataLoader(train_set, batch_size=batch_size, sampler=sampler, num_workers=num_workers,
                       pin_memory=pin_memory)
    test = DataLoader(test_set, batch_size=128, shuffle=False, num_workers=num_workers, pin_memory=pin_memory)
    return train, test


class DevicePrefetcher:
    """Wraps an iterable of (inputs, labels) host batches: copies the next ``depth`` batches to the device on a side
    stream from pinned memory while the current step computes, and yields device tensors that are safe to use on the
    current stream.

    The device side is a fixed ring of ``depth + 1`` buffers owned by the prefetcher - no allocation per step (a
    caching-allocator miss means ``cudaMalloc``, which on a busy or virtualised host costs more than the copy it
    serves). A buffer is overwritten only after the compute stream has passed the work that was enqueued while its
    batch was current (an event recorded when the consumer asks for the next batch), so the yielded tensors are valid
    until ``depth`` further batches have been requested. Batches whose shape differs from the ring's (ragged last
    batch) get tensors of their own."""

    def __init__(self, loader, device, depth=2):
        self.loader, self.device = loader, torch.device(device)
        self.depth = max(1, int(depth))
        self.stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None
        self._ring = []

    def __len__(self):
        return len(self.loader)

    def _slot_for(self, x, y, index):
        if len(self._ring) <= index:
            self._ring.append({"x": torch.empty(x.shape, dtype=x.dtype, device=self.device),
                               "y": torch.empty(y.shape, dtype=y.dtype, device=self.device), "free": None})
        slot = self._ring[index]
        if slot["x"].shape != x.shape or slot["x"].dtype != x.dtype or slot["y"].shape != y.shape \
                or slot["y"].dtype != y.dtype:
            return None
        return slot

    def _stage(self, batch, index):
        x, y = batch[0], batch[1]
        if self.stream is None:
            return {"xd": x.to(self.device), "yd": y.to(self.device), "ev": None, "slot": None, "host": None}
        if not x.is_pinned():
            x = x.pin_memory()
        if not y.is_pinned():
            y = y.pin_memory()
        slot = self._slot_for(x, y, index)
        with torch.cuda.stream(self.stream):
            if slot is not None:
                if slot["free"] is not None:
                    self.stream.wait_event(slot["free"])      # its previous batch has been consumed on the device
                xd, yd = slot["x"], slot["y"]
                xd.copy_(x, non_blocking=True)
                yd.copy_(y, non_blocking=True)
            else:
                xd = x.to(self.device, non_blocking=True)
                yd = y.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)      # this batch's copies only: later copies must not hold the consumer back
        return {"xd": xd, "yd": yd, "ev": ev, "slot": slot, "host": (x, y)}

    def __iter__(self):
        import collections
        it = iter(self.loader)
        staged = collections.deque()
        nbuf = self.depth + 1
        counter = [0]

        def fill():
            while len(staged) < self.depth:
                try:
                    batch = next(it)
                except StopIteration:
                    return
                staged.append(self._stage(batch, counter[0] % nbuf))
                counter[0] += 1

        fill()
        while staged:
            cur = staged.popleft()
            fill()
            if self.stream is not None:
                cs = torch.cuda.current_stream(self.device)
                cs.wait_event(cur["ev"])
                if cur["slot"] is None:
                    cur["xd"].record_stream(cs)
                    cur["yd"].record_stream(cs)
            yield cur["xd"], cur["yd"]
            # the consumer is back for the next batch: everything it enqueued for this one is on its stream now
            if self.stream is not None and cur["slot"] is not None:
                free = torch.cuda.Event()
                free.record(torch.cuda.current_stream(self.device))
                cur["slot"]["free"] = free
