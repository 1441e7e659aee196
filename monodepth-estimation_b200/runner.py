"""Batched replacement for the data path of ``run_mars_depth.py:run_depth_estimation`` (``run_mars_depth.py:78-147``).

The reference's loop is strictly one image at a time — open, LANCZOS resize, ``float()/255``, blocking ``.to(device)``,
``model(x, metadata)``, blocking ``.cpu()`` — and its ``--batch_size`` flag is parsed and never read (``:47-50``).  Once a
forward costs ~70 us per image that loop is all host work, so this module does the same steps as a pipeline:

    decode pool (host threads: PIL decode + LANCZOS, both release the GIL)  ->  shape/source buckets of ``batch_size`` images
      ->  pinned uint8 [B,H,W,3] batches  ->  ``BatchPipeline`` (H2D | forward | D2H on three streams)  ->  depth = 1 / disp.clamp(1e-6)

Per image the arithmetic is the reference's: the same ``Image.open(...).convert("RGB")``, the same ``int(size * scale)`` LANCZOS
target (``:94-100``), the same path-based source detection (``:106-117``), ``float()/255`` evaluated on the device bit-identically
(``mde_u8hwc_to_f32chw``), the same ``1.0 / disp.clamp(min=1e-6)`` (``:137``), and ``<name>/<name>_depth.npy`` written like ``:190``.
Images are independent in the eval forward, so batching changes no value (``tests/test_forward_gpu.py::test_batch_shard_invariance``)
— with one stated exception inherited from the model: the input range heuristic (``encoder.py:132-135``) looks at the whole batch,
so images that would take DIFFERENT branches of it on their own (not possible for 8-bit images with the stock ``normalize``
statistics) must not share a batch; ``batch_size=1`` reproduces the reference loop exactly in every case.
Visualisation (matplotlib / plotly figures, ``:150-260``) is out of scope.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch


def detect_source(image_path: str, source: str = "auto") -> str:
    """run_mars_depth.py:106-117."""
    if source != "auto":
        return source
    p = image_path.lower()
    if "rover" in p:
        return "rover"
    if "ingenuity" in p or "aerial" in p:
        return "ingenuity"
    if "satellite" in p or "orbital" in p:
        return "satellite"
    return "rover"


def load_image_u8(image_path: str, max_size: int = 1024) -> np.ndarray:
    """run_mars_depth.py:94-100: decode to RGB, LANCZOS-resize so that the longer side is at most ``max_size`` -> uint8 [H,W,3]."""
    from PIL import Image

    img = Image.open(image_path).convert("RGB")
    if max(img.size) > max_size:
        scale = max_size / max(img.size)
        img = img.resize((int(img.size[0] * scale), int(img.size[1] * scale)), Image.LANCZOS)
    return np.ascontiguousarray(np.array(img))


def iter_batches(image_paths, batch_size: int = 1, max_size: int = 1024, source: str = "auto", workers: int | None = None, prefetch: int | None = None):
    """Decode on a thread pool and yield ``(indices, uint8 [B,H,W,3] array, source type)`` batches: images are bucketed by
    (height, width, source type) in arrival order, a bucket is emitted when it holds ``batch_size`` images, the rest at the end.
    ``prefetch`` bounds the number of decoded-but-unconsumed images (default 4 batches)."""
    if batch_size < 1:
        raise ValueError("batch_size must be >= 1")
    paths = list(image_paths)
    workers = workers or min(32, (os.cpu_count() or 4))
    prefetch = prefetch or max(4 * batch_size, workers)
    buckets: "OrderedDict[tuple, list]" = OrderedDict()
    with ThreadPoolExecutor(max_workers=workers) as pool:
        pending = []
        nxt = 0

        def top_up():
            nonlocal nxt
            while nxt < len(paths) and len(pending) < prefetch:
                pending.append((nxt, pool.submit(load_image_u8, paths[nxt], max_size)))
                nxt += 1

        top_up()
        while pending:
            idx, fut = pending.pop(0)
            arr = fut.result()  # in submission order: results and buckets are deterministic
            top_up()
            key = (arr.shape[0], arr.shape[1], detect_source(paths[idx], source))
            b = buckets.setdefault(key, [])
            b.append((idx, arr))
            if len(b) == batch_size:
                del buckets[key]
                yield [i for i, _ in b], np.stack([a for _, a in b]), key[2]
    for key, b in buckets.items():
        yield [i for i, _ in b], np.stack([a for _, a in b]), key[2]


def run_depth_estimation(model, image_paths, output_dir: str | None = None, batch_size: int = 1, max_size: int = 1024, source: str = "auto",
                         workers: int | None = None, keep_results: bool = True):
    """The reference's ``run_depth_estimation`` data path with ``batch_size`` honoured.  Returns one dict per input image, in input
    order: ``{"path", "source", "disp", "depth", "uncertainty"}`` (NumPy float32 [H,W]; uncertainty None for the standard model),
    and writes ``<output_dir>/<name>/<name>_depth.npy`` when ``output_dir`` is given.  ``keep_results=False`` keeps only the paths
    (large runs: the maps are on disk)."""
    from .pipeline import BatchPipeline

    paths = list(image_paths)
    dev = next(model.parameters()).device
    if dev.type != "cuda":
        raise RuntimeError("the model must live on a CUDA (sm_100) device: there is no CPU path — call model.to('cuda')")
    enhanced = hasattr(model, "uncertainty_module")
    keys = (("disp", 0),) + ((("uncertainty", 0),) if enhanced else ())
    results = [None] * len(paths)
    if output_dir is not None:
        os.makedirs(output_dir, exist_ok=True)
    pipes = {}  # one pipeline per (batch shape, source): its ring buffers are shape-specific

    def consume_for(indices, src):
        def consume(_, host):
            disp_b = host[("disp", 0)].numpy()
            unc_b = host[("uncertainty", 0)].numpy() if enhanced else None
            for j, i in enumerate(indices):
                disp = disp_b[j, 0].copy()
                depth = np.float32(1.0) / np.maximum(disp, np.float32(1e-6))  # :137, the same IEEE division the device would do
                rec = {"path": paths[i], "source": src, "disp": disp, "depth": depth, "uncertainty": None if unc_b is None else unc_b[j, 0].copy()}
                if output_dir is not None:
                    name = os.path.splitext(os.path.basename(paths[i]))[0]
                    d = os.path.join(output_dir, name)
                    os.makedirs(d, exist_ok=True)
                    np.save(os.path.join(d, f"{name}_depth.npy"), depth)  # :190
                if not keep_results:
                    rec = {"path": paths[i], "source": src}
                results[i] = rec
        return consume

    it = iter(iter_batches(paths, batch_size, max_size, source, workers))
    carry = next(it, None)
    while carry is not None:
        indices0, batch0, src = carry
        key = (tuple(batch0.shape), src)
        pipe = pipes.get(key)
        if pipe is None:
            if len(pipes) >= 4:
                pipes.pop(next(iter(pipes)))
            pipe = pipes[key] = BatchPipeline(model, keys=keys, depth=3)
        index_lists = []
        carry = None

        def same_key_run(first=(indices0, batch0)):
            """Consecutive batches of one (shape, source): ONE pipe.run call, so H2D(i+1), forward(i) and D2H(i-1) overlap."""
            nonlocal carry
            item = first
            while True:
                index_lists.append(item[0])
                yield torch.from_numpy(item[1]).pin_memory()
                nxt = next(it, None)
                if nxt is None:
                    return
                if (tuple(nxt[1].shape), nxt[2]) != key:
                    carry = nxt
                    return
                item = (nxt[0], nxt[1])

        consumers = {}

        def consume(bi, host, _src=src):
            consumers.setdefault(bi, consume_for(index_lists[bi], _src))(bi, host)

        pipe.run(same_key_run(), {"source": src}, consume)
    return results


def main(argv=None):
    """``python -m monodepth_estimation_b200.runner --input DIR_OR_FILE [--batch_size N] ...``: the subset of run_mars_depth.py's
    flags that concerns the data path (``--input --output_dir --source --max_size --batch_size --enhanced_model``)."""
    import argparse

    from . import DepthAnything, EnhancedDepthAnything

    ap = argparse.ArgumentParser(description="Mars terrain depth estimation (batched data path)")
    ap.add_argument("--input", required=True)
    ap.add_argument("--output_dir", default="./outputs")
    ap.add_argument("--source", default="auto", choices=["auto", "rover", "ingenuity", "satellite"])
    ap.add_argument("--max_size", type=int, default=1024)
    ap.add_argument("--batch_size", type=int, default=1)
    ap.add_argument("--enhanced_model", action="store_true")
    ap.add_argument("--weights", default=None, help="state_dict file (torch.save); default: random init, as the reference without a checkpoint")
    a = ap.parse_args(argv)
    exts = (".png", ".jpg", ".jpeg", ".tif", ".tiff", ".bmp")
    if os.path.isdir(a.input):
        paths = sorted(os.path.join(a.input, f) for f in os.listdir(a.input) if f.lower().endswith(exts))
    else:
        paths = [a.input]
    model = (EnhancedDepthAnything if a.enhanced_model else DepthAnything)(pretrained=False)
    if a.weights:
        model.load_state_dict(torch.load(a.weights, map_location="cpu"))
    model = model.eval().to("cuda")
    res = run_depth_estimation(model, paths, a.output_dir, a.batch_size, a.max_size, a.source, keep_results=False)
    print(f"processed {len(res)} image(s) -> {a.output_dir}")


if __name__ == "__main__":
    main()
