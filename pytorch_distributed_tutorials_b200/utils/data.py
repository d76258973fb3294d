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
    """Wraps an iterable of (inputs, labels) host batches: copies batch i+1 to the device on a side stream from
    pinned memory while step i computes. Yields device tensors already safe to use on the current stream."""

    def __init__(self, loader, device):
        self.loader, self.device = loader, torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None

    def __len__(self):
        return len(self.loader)

    def _stage(self, batch):
        x, y = batch[0], batch[1]
        if self.stream is None:
            return x.to(self.device), y.to(self.device)
        if not x.is_pinned():
            x = x.pin_memory()
        if not y.is_pinned():
            y = y.pin_memory()
        with torch.cuda.stream(self.stream):
            xd = x.to(self.device, non_blocking=True)
            yd = y.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)      # this batch's copies only: later copies must not hold the consumer back
        return xd, yd, ev, (x, y)

    def __iter__(self):
        it = iter(self.loader)
        try:
            nxt = self._stage(next(it))
        except StopIteration:
            return
        while nxt is not None:
            cur = nxt
            try:
                nxt = self._stage(next(it))
            except StopIteration:
                nxt = None
            if self.stream is not None:
                torch.cuda.current_stream(self.device).wait_event(cur[2])
                cur[0].record_stream(torch.cuda.current_stream(self.device))
                cur[1].record_stream(torch.cuda.current_stream(self.device))
            yield cur[0], cur[1]
