"""Collate functions and loaders of the reference's data layer.

  video_collate_fn         vivit_transformer/vivit_classifier/data_config/dataloader.py:6-50
                           -> {'pixel_values': list of (T,224,224,3) uint8, 'labels': LongTensor(B)}
  clip_collate_fn          videoswintransformer/swin_video_classifier/data_config/dataloader.py:5-19
                           -> (clips (B,n,C,T,H,W), labels (B,n))
  collate_fn               model_configuration.py:11-31 (torch.stack of per-example pixel_values)
  BatchLoader / create_dataloaders   :52-134 -- same arguments; batches are built on the GPU with one
                           fused launch per batch (``dataset.get_batch``), no worker processes: CUDA
                           contexts do not survive DataLoader fork workers.
"""
import numpy as np
import torch


def video_collate_fn(batch):
    pixel_values, labels = [], []
    for sample in batch:
        pv = sample.get("pixel_values")
        if pv is not None:
            if pv.ndim != 4 or pv.shape[-1] != 3:
                pv = pv.squeeze()
            if pv.ndim != 4 or pv.shape[-1] != 3:
                raise ValueError(f"unexpected frame shape in collate_fn: {tuple(sample['pixel_values'].shape)}")
            pixel_values.append(pv)
        if sample.get("labels") is not None:
            labels.append(torch.as_tensor(sample["labels"]))
    labels = torch.stack(labels) if labels else torch.zeros(len(batch), dtype=torch.long)
    return {"pixel_values": pixel_values, "labels": labels}


def clip_collate_fn(batch):
    min_clips = min(c.size(0) for c, _ in batch)
    clips = torch.stack([c[:min_clips] for c, _ in batch], dim=0)
    labels = torch.stack([l[:min_clips] for _, l in batch], dim=0)
    return clips, labels


def collate_fn(examples):
    pixel_values = torch.stack([torch.as_tensor(e["pixel_values"]) for e in examples])
    labels = torch.tensor([int(e["labels"]) for e in examples])
    return {"pixel_values": pixel_values, "labels": labels}


class BatchLoader:
    """Iterates a dataset in batches built by ``dataset.get_batch`` (one fused kernel launch each).
    ``rank``/``world_size`` shard the videos v -> rank (v mod world) with no collective."""

    def __init__(self, dataset, batch_size=4, shuffle=False, drop_last=False, seed=42, rank=0, world_size=1):
        self.dataset, self.batch_size, self.shuffle, self.drop_last = dataset, int(batch_size), shuffle, drop_last
        self.rank, self.world_size = rank, world_size
        self._gen = np.random.default_rng(seed)

    def order(self):
        idx = np.arange(len(self.dataset))
        if self.shuffle:
            idx = self._gen.permutation(len(self.dataset))
        return [int(i) for i in idx[self.rank::self.world_size]]

    def __len__(self):
        n = len(range(self.rank, len(self.dataset), self.world_size))
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    def __iter__(self):
        order = self.order()
        for s in range(0, len(order), self.batch_size):
            chunk = order[s:s + self.batch_size]
            if self.drop_last and len(chunk) < self.batch_size:
                return
            yield self.dataset.get_batch(chunk)


def create_dataloaders(args, sampling_methods, logger=None, frame_sources=None, dataset_cls=None, rank=0, world_size=1):
    """``create_dataloaders(args, sampling_methods, logger)`` of the reference: ``args`` needs data_dir,
    num_frames, batch_size (and optionally test_data_dir); returns (train, val, test) loaders."""
    from .datasets import VivitVideoDataset

    dataset_cls = dataset_cls or VivitVideoDataset
    loaders = []
    for split in ("train", "val", "test"):
        root = getattr(args, "test_data_dir", None) if split == "test" and getattr(args, "test_data_dir", None) else args.data_dir
        src = None if frame_sources is None else frame_sources[split]
        ds = dataset_cls(root, mode=split, sampling_method=sampling_methods[split], num_frames=args.num_frames,
                         logger=logger, frame_source=src)
        loaders.append(BatchLoader(ds, args.batch_size, shuffle=(split == "train"), rank=rank, world_size=world_size))
    return tuple(loaders)
