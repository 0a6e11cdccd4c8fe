"""The drop-in Dataset / collate / image-processor / transform layer on the GPU, driven the way the
reference's trainers drive theirs, against the oracle."""
import random
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import processors as op
from oracle import samplers as osm

pytestmark = pytest.mark.gpu


def _source(n=3, frames=40, h=180, w=320):
    from lvc_b200.datasets import SyntheticFrameSource
    from lvc_b200.synth import synth_video

    return SyntheticFrameSource(n, frames, h, w), [synth_video(1000 + i, frames, h, w) for i in range(n)]


def test_vivit_dataset_collate_processor_like_process_batch():
    from lvc_b200.dataloader import video_collate_fn
    from lvc_b200.datasets import VideoDataset
    from lvc_b200.processors import VivitImageProcessor

    src, host = _source()
    ds = VideoDataset(None, mode="val", sampling_method="random_window", num_frames=16, frame_source=src)
    assert len(ds) == 3 and ds.class_labels == ["non_referral", "referral"] and len(ds.video_paths) == 3
    samples = [ds[i] for i in range(3)]
    rng = random.Random(42)  # the reference seeds once in __init__, then one global stream in access order
    for i, s in enumerate(samples):
        want_idx = osm.dataset_indices("random_window", 40, 16, rng)
        assert s["frame_indices"] == want_idx and s["video_path"] == ds.video_paths[i]
        assert s["pixel_values"].dtype == torch.uint8 and tuple(s["pixel_values"].shape) == (16, 224, 224, 3)
        assert np.array_equal(s["pixel_values"].cpu().numpy(), op.resize_clip_cv2(host[i][want_idx]))
        assert int(s["labels"]) == ds.labels[i]
    batch = video_collate_fn(samples)
    assert isinstance(batch["pixel_values"], list) and batch["labels"].tolist() == ds.labels
    # vivit .../trainers/trainer.py:22-26,62-95
    proc = VivitImageProcessor(num_frames=16, image_size=224, patch_size=16)
    inputs = proc(batch["pixel_values"], return_tensors="pt")
    assert set(inputs.keys()) == {"pixel_values"}
    inputs = {k: v.to("cuda") for k, v in inputs.items()}
    pv = inputs["pixel_values"]
    assert tuple(pv.shape) == (3, 16, 3, 224, 224) and pv.dtype == torch.float32
    for i, s in enumerate(samples):
        assert np.array_equal(pv[i].cpu().numpy(), op.vivit_processor(s["pixel_values"].cpu().numpy()))
    # per-video call with a list of HWC frames, numpy input (trainer.py:75-78)
    one = proc(list(samples[0]["pixel_values"].cpu().numpy()), return_tensors="pt")["pixel_values"]
    assert tuple(one.shape) == (1, 16, 3, 224, 224) and torch.equal(one[0], pv[0])


def test_timesformer_processor_call_pattern():
    from lvc_b200.processors import TimesformerImageProcessor

    fr = np.random.default_rng(0).integers(0, 256, (8, 224, 224, 3), dtype=np.uint8)
    proc = TimesformerImageProcessor(image_mean=[0.5] * 3, image_std=[0.5] * 3)
    proc.size = {"height": 224, "width": 224}
    proc.crop_size = {"height": 224, "width": 224}
    out = proc(images=list(fr), return_tensors="pt", do_resize=True, size={"height": 224, "width": 224},
               do_center_crop=True, crop_size={"height": 224, "width": 224})["pixel_values"]
    assert tuple(out.shape) == (1, 8, 3, 224, 224)
    assert np.array_equal(out[0].cpu().numpy(), op.videomae_processor(fr, (0.5,) * 3, (0.5,) * 3))
    with pytest.raises(NotImplementedError):
        proc(images=list(fr), size={"height": 256, "width": 256})
    with pytest.raises(ValueError):
        proc(images=list(fr.astype(np.float32)))


def test_swin_and_lstm_datasets_follow_the_pytorchvideo_path():
    from lvc_b200.dataloader import clip_collate_fn
    from lvc_b200.datasets import LstmVideoDataset, SwinVideoDataset, window_subsample

    src, host = _source(n=2, frames=50, h=240, w=320)
    ds = SwinVideoDataset(None, mode="val", sampling_method="random", num_frames=8, frame_source=src)
    idx, dyn = ds.get_sampling_indices(ds.video_paths[0], 50)
    assert dyn is None and idx == osm.dataset_indices("random", 50, 8, random.Random(42))
    _, dyn = ds.get_sampling_indices(ds.video_paths[0], 5)
    assert dyn == osm.dynamic_fps(5, 8, 30.0)
    ds = SwinVideoDataset(None, mode="val", sampling_method="random", num_frames=8, frame_source=src)
    rng = random.Random(42)
    items = [ds[i] for i in range(2)]
    for i, (clip, label) in enumerate(items):
        raw = osm.dataset_indices("random", 50, 8, rng)
        eff = [min(raw) + k for k in osm.uniform_temporal_subsample(max(raw) - min(raw) + 1, 8)]
        assert window_subsample(raw, 8) == eff
        assert tuple(clip.shape) == (1, 3, 8, 224, 224) and tuple(label.shape) == (1,)
        assert np.allclose(clip[0].cpu().numpy(), op.ptv_eval_chain(host[i][eff]), atol=1.1e-2, rtol=1e-5)
    clips, labels = clip_collate_fn(items)
    assert tuple(clips.shape) == (2, 1, 3, 8, 224, 224) and tuple(labels.shape) == (2, 1)
    lstm = LstmVideoDataset(None, split="val", sampling_method="random_window", sequence_length=8, frame_source=src)
    a, la = lstm[0]
    b, _ = lstm[0]  # re-seeded with 42 on every call: identical draws
    assert torch.equal(a, b) and tuple(a.shape) == (3, 8, 224, 224) and la.dtype == torch.float32


def test_ptv_transform_and_sharded_loader():
    from lvc_b200.dataloader import BatchLoader, create_dataloaders
    from lvc_b200.datasets import VideoDataset
    from lvc_b200.processors import PtvEvalTransform

    src, host = _source(n=5, frames=20, h=120, w=200)
    v = torch.from_numpy(host[0]).permute(3, 0, 1, 2).float()  # what EncodedVideo.get_clip returns
    out = PtvEvalTransform(8)({"video": v, "label": 1})
    eff = osm.uniform_temporal_subsample(20, 8)
    assert out["label"] == 1 and np.allclose(out["video"].cpu().numpy(), op.ptv_eval_chain(host[0][eff]), atol=1.1e-2, rtol=1e-5)
    ds = VideoDataset(None, sampling_method="uniform", num_frames=4, frame_source=src)
    seen = []
    for rank in range(2):
        for b in BatchLoader(ds, batch_size=2, rank=rank, world_size=2):
            assert len(b["pixel_values"]) == len(b["video_paths"]) <= 2
            seen += b["video_paths"]
    assert sorted(seen) == sorted(ds.video_paths)
    args = SimpleNamespace(data_dir=None, num_frames=4, batch_size=2)
    tr, va, te = create_dataloaders(args, {"train": "random", "val": "uniform", "test": "uniform"}, None,
                                    frame_sources={"train": src, "val": src, "test": src})
    assert len(va) == 3 and sum(len(b["pixel_values"]) for b in va) == 5


def test_clip_embed_pipeline_graph_replay_matches_eager():
    """CUDA-graph pipeline (table upload -> clip kernel -> GEMM) == the two eager calls, across changing index tables,
    ROI boxes and swapped videos."""
    import numpy as np
    import torch

    from lvc_b200 import ClipBuilder, sampling
    from lvc_b200.patch_embed import TimesformerPatchEmbeddings
    from lvc_b200.pipeline import ClipEmbedPipeline
    from lvc_b200.synth import synth_video

    T = 4
    vids = [torch.from_numpy(synth_video(40 + i, 12, 240, 320)).cuda() for i in range(3)]
    cb = ClipBuilder("timesformer", num_frames=T, layout="TOKENS", patch=(1, 16, 16))
    m = TimesformerPatchEmbeddings(out_dtype=torch.float32).cuda()
    pipe = ClipEmbedPipeline(cb, m, vids, T, boxes=True)
    rng = sampling.SamplerRng(3)
    g = np.random.default_rng(0)
    for it in range(4):
        idx = np.array([sampling.get_sampling_indices("random", 12, T, rng) for _ in vids], np.int32)
        boxes = np.full((3, T, 4), -1, np.int32)
        if it % 2:
            x1, y1 = g.integers(0, 100, (3, T)), g.integers(0, 80, (3, T))
            boxes = np.stack([x1, y1, x1 + 150, y1 + 120], -1).astype(np.int32)
        if it == 2:
            vids = [torch.from_numpy(synth_video(50 + i, 12, 240, 320)).cuda() for i in range(3)]
            pipe.set_videos(vids)
        got = pipe(idx, boxes).clone()
        want = m.forward_tokens(cb(vids, idx, boxes=boxes), 3, T)
        assert torch.equal(got, want)
    assert pipe.replays == 4
    with pytest.raises(ValueError):
        pipe.set_videos([torch.zeros(2, 100, 100, 3, dtype=torch.uint8, device="cuda")] * 3)


def test_pipeline_back_to_back_calls_while_the_gpu_is_behind():
    """the host rewrote the single pinned table for call k+1 while replay k's own upload node had not run
    yet (GPU busy), so batch k was built with batch k+1's indices.  Two graphs with private tables alternate now: with
    the device held back, back-to-back calls with different index tables must each produce their own batch."""
    import numpy as np
    import torch

    from lvc_b200 import ClipBuilder
    from lvc_b200.patch_embed import TimesformerPatchEmbeddings
    from lvc_b200.pipeline import ClipEmbedPipeline
    from lvc_b200.synth import synth_video

    T = 4
    vids = [torch.from_numpy(synth_video(60 + i, 16, 120, 160)).cuda() for i in range(2)]
    cb = ClipBuilder("timesformer", num_frames=T, layout="TOKENS", patch=(1, 16, 16))
    m = TimesformerPatchEmbeddings(out_dtype=torch.float32).cuda()
    pipe = ClipEmbedPipeline(cb, m, vids, T)
    tables = [np.array([[k, k + 1, k + 2, k + 3], [15 - k, 14 - k, 13 - k, 12 - k]], np.int32) for k in range(6)]
    want = [m.forward_tokens(cb(vids, t), 2, T).clone() for t in tables]
    torch.cuda.synchronize()
    got = []
    torch.cuda._sleep(int(2e8))            # ~0.1 s of device time: every replay below is enqueued behind it
    for t in tables:
        got.append(pipe(t).clone())        # clone is stream-ordered: taken before the slot's buffer is reused
    torch.cuda.synchronize()
    for k in range(6):
        assert torch.equal(got[k], want[k]), k
