"""The batched hot path as one object: scored candidate boxes of a batch of shelf images -> NMS -> crops -> encoder
(torch conv trunk + MAC/R-MAC kernel) -> cosine top-k against a prepared gallery.

This is what bench.py times and what a production caller uses when it has more than one image in flight; it performs ONE
host synchronisation per step (the kept-box counts, needed to size the crop tensor) instead of the reference's four per
box (cvpce/production.py:20) plus one per result element (:70).
"""
import torch

from . import classification, ops


class StageTimer:
    """Records CUDA events around each library call so that bench.py can report per-kernel durations measured inside
    the timed region (events are recorded on the current stream, where the kernels are launched)."""

    def __init__(self):
        self.records = []  # (name, start_event, end_event, algorithmic bytes, flops)

    def span(self, name, nbytes=0.0, flops=0.0):
        timer = self

        class _Span:
            def __enter__(self):
                self.s = torch.cuda.Event(enable_timing=True)
                self.e = torch.cuda.Event(enable_timing=True)
                self.s.record()

            def __exit__(self, *exc):
                self.e.record()
                timer.records.append((name, self.s, self.e, float(nbytes), float(flops)))

        return _Span()

    def summary(self):
        out = {}
        for name, s, e, nbytes, flops in self.records:
            d = out.setdefault(name, {"launches": 0, "ms": 0.0, "bytes": 0.0, "flops": 0.0})
            d["launches"] += 1
            d["ms"] += s.elapsed_time(e)
            d["bytes"] += nbytes
            d["flops"] += flops
        return out


class _NoTimer:
    def span(self, *a, **k):
        import contextlib
        return contextlib.nullcontext()


class _Uploads(list):
    """(event, first image, end image) per upload group, plus the staging buffer they went to."""

    def __init__(self, slot):
        super().__init__()
        self.slot = slot


class HotPath:
    def __init__(self, encoder, gallery_embedding, k=1, size=256, antialias=True, iou_threshold=0.5, detections_per_img=200,
                 min_score=None, trunk_batch=128, trunk_dtype=torch.float32, channels_last=False, exact=True, rmac_levels=0,
                 n_streams=1, pad_channel=True):
        self.encoder = encoder
        self.pad_channel = pad_channel
        self.gallery = classification.PreparedGallery(gallery_embedding)
        self.k = k
        self.size = size
        self.antialias = antialias
        self.iou_threshold = iou_threshold
        self.detections_per_img = detections_per_img
        self.min_score = min_score
        self.trunk_batch = trunk_batch
        self.trunk_dtype = trunk_dtype
        self.channels_last = channels_last
        self.exact = exact
        self.rmac_levels = rmac_levels
        self.n_streams = n_streams
        self._streams = None
        self._copy_stream = None
        self._stage, self._stage_key, self._stage_free, self._stage_next = [None, None], [None, None], [None, None], 0
        self.affine = classification.tanh_normalize_affine()  # scale_to_tanh + MACVGG normalisation folded into the crop kernel
        self.launches = 0  # launches of OUR kernels (library calls), counted per call site
        self.timer = _NoTimer()

    def _side_streams(self, dev):
        if self._streams is None:
            self._streams = [torch.cuda.Stream(device=dev) for _ in range(self.n_streams)]
        return self._streams

    @torch.no_grad()
    def step(self, images, cand_boxes, cand_scores, seg_offsets, max_n, uploads=None):
        """images: list of CUDA f32 [3,H,W]; cand_boxes f32[total,4] / cand_scores f32[total] / seg_offsets i32[B+1] on the
        device (original-image pixels); max_n: host upper bound of candidates per image.
        Returns dict(boxes f32[K,4], image_idx i32[K], counts i32[B] (host), topk_idx i64[K,k], topk_sim f32[K,k], embedding)."""
        B = len(images)
        total = cand_boxes.shape[0]
        with self.timer.span("nms", 20.0 * total):
            keep, counts, kept_boxes, _ = ops.nms_batched(cand_boxes, cand_scores, seg_offsets, max_n, self.iou_threshold,
                                                          self.min_score, self.detections_per_img, want_boxes=True)
        self.launches += 1
        return self._classify_kept(images, counts, kept_boxes, seg_offsets[:-1].to(torch.int64), uploads)

    @torch.no_grad()
    def step_from_head(self, images, cls_logits, bbox_regression, anchors, frame_sizes, box_scale, score_thresh=0.05,
                       topk_candidates=1000, uploads=None):
        """The same step starting one stage earlier: the detector head's per-level outputs of the batch
        (tv:retinanet.py:508-571) -> candidate selection -> NMS over its fixed-stride segments, boxes rescaled to the
        original images by box_scale f32[B,2] (tv:transform.py:306-319) -> crops -> ... as in step()."""
        B = len(images)
        n_anchors = sum(int(l.shape[1]) for l in cls_logits)
        with self.timer.span("select_candidates", 4.0 * n_anchors * B):
            cb, cs, _, seg_counts, _, stride = ops.select_candidates(cls_logits, bbox_regression, anchors, frame_sizes,
                                                                     score_thresh, topk_candidates, want_labels=False)
        with self.timer.span("nms", 20.0 * B * stride):
            _, counts, kept_boxes, _ = ops.nms_segments(cb, cs, stride, seg_counts, self.iou_threshold, self.min_score,
                                                        self.detections_per_img, box_scale=box_scale, want_boxes=True)
        self.launches += 2
        return self._classify_kept(images, counts, kept_boxes, torch.arange(B, device=cb.device, dtype=torch.int64) * stride, uploads)

    def upload_images(self, images_h, upload_group=4):
        """Starts the host->device transfer of pinned host images on a copy stream, in groups of `upload_group`, into a
        persistent staging buffer.  Returns (device images, uploads) for step(..., uploads=uploads) /
        step_from_head(..., uploads=uploads): the step then cuts the crops of a group as soon as its images have landed
        and runs the trunk on the crops that exist, so the transfer of the step's own images is hidden behind its
        convolutions instead of preceding them.
        There are TWO staging buffers, used alternately: a serving loop may call upload_images for step s+1 before it runs
        step s, and the copy stream then moves the next step's images while this step computes (the buffer a call takes is
        the one released two steps ago)."""
        dev = self.gallery.f32.device
        B = len(images_h)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        slot = self._stage_next
        self._stage_next ^= 1
        shapes = [tuple(im.shape) for im in images_h]
        if self._stage[slot] is None or self._stage_key[slot] != shapes:
            self._stage[slot] = [torch.empty(sh, dtype=torch.float32, device=dev) for sh in shapes]
            self._stage_key[slot] = shapes
        stage = self._stage[slot]
        cs = self._copy_stream
        if self._stage_free[slot] is not None:
            cs.wait_event(self._stage_free[slot])  # the step that last used this buffer has finished reading it
            self._stage_free[slot] = None
        uploads = _Uploads(slot)
        with torch.cuda.stream(cs):
            for lo in range(0, B, upload_group):
                hi = min(B, lo + upload_group)
                for i in range(lo, hi):
                    stage[i].copy_(images_h[i], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                uploads.append((ev, lo, hi))
        return stage, uploads

    @torch.no_grad()
    def step_host(self, images_h, cand_boxes_h, cand_scores_h, seg_offsets_h, max_n, upload_group=4):
        """step() for inputs in PINNED HOST memory (see upload_images).  Same results as step()."""
        dev = self.gallery.f32.device
        images, uploads = self.upload_images(images_h, upload_group)
        return self.step(images, cand_boxes_h.to(dev, non_blocking=True), cand_scores_h.to(dev, non_blocking=True),
                         seg_offsets_h.to(dev, non_blocking=True), max_n, uploads=uploads)

    def _classify_kept(self, images, counts, kept_boxes, seg_begin, uploads=None):
        B = len(images)
        counts_h = ops.checked_counts(counts.cpu(), "HotPath.step")  # the one host synchronisation of the step
        K = int(counts_h.sum())
        dev = kept_boxes.device
        counts64 = counts.to(torch.int64)
        image_idx = torch.repeat_interleave(torch.arange(B, device=dev), counts64, output_size=K)
        first = torch.cumsum(counts64, 0) - counts64
        pos = seg_begin[image_idx] + (torch.arange(K, device=dev) - first[image_idx])
        boxes = kept_boxes[pos]
        table = ops.make_image_table(images)
        S = self.size
        crop_bytes = 12.0 * S * S * K + 12.0 * float(((boxes[:, 2].long() - boxes[:, 0].long()).clamp(min=0) *
                                                      (boxes[:, 3].long() - boxes[:, 1].long()).clamp(min=0)).sum()) \
            if not isinstance(self.timer, _NoTimer) else 0.0
        image_idx32 = image_idx.to(torch.int32)
        # fast trunk: 4-channel crops (zero 4th channel) so that cuDNN skips its channel-padding pass before conv1_1
        # (fp32 / TF32 trunk only: with half-precision activations the 4-channel first layer measured slower, 25.3k -> 24.3k crops/s)
        pad4 = bool(self.channels_last and self.pad_channel and self.trunk_dtype == torch.float32
                    and getattr(self.encoder, "fast_trunk", False) and getattr(self.encoder, "accepts_padded_input", False))
        if uploads is None:
            with self.timer.span("crop_resize", crop_bytes):
                crops = ops.crop_resize(table, B, boxes, image_idx32, S, self.antialias, 0.5, self.affine,
                                        self.trunk_dtype, self.channels_last, pad_channel=pad4)
            self.launches += 1
            ensure_crops = None
        else:
            # images arrive group by group (step_host): one crop launch per group, issued right before the first trunk batch
            # that needs it, after the main stream has waited for the group's upload
            phys = torch.empty((K, S, S, 4 if pad4 else 3) if self.channels_last else (K, 3, S, S), dtype=self.trunk_dtype, device=dev)
            crops = phys.permute(0, 3, 1, 2) if self.channels_last else phys
            first_crop = [0]
            for c in counts_h.tolist():
                first_crop.append(first_crop[-1] + int(c))
            pending = list(uploads)
            main_stream = torch.cuda.current_stream(dev)

            def ensure_crops(upto):
                while pending and first_crop[pending[0][1]] < upto:
                    ev, lo, hi = pending.pop(0)
                    a, b = first_crop[lo], first_crop[hi]
                    main_stream.wait_event(ev)
                    if b > a:
                        with self.timer.span("crop_resize", crop_bytes * (b - a) / max(K, 1)):
                            ops.crop_resize(table, B, boxes[a:b], image_idx32[a:b], S, self.antialias, 0.5, self.affine,
                                            self.trunk_dtype, self.channels_last, out=phys[a:b], pad_channel=pad4)
                        self.launches += 1
                slot = getattr(uploads, "slot", 0)
                if not pending and self._stage_free[slot] is None:
                    self._stage_free[slot] = torch.cuda.Event()
                    self._stage_free[slot].record(main_stream)
        D = self.encoder.embedding_size
        emb = torch.empty((K, D), dtype=torch.float32, device=dev)
        regions = None
        own0 = getattr(self.encoder, "own_launches", 0)
        if hasattr(self.encoder, "stage_timer"):
            self.encoder.stage_timer = None if isinstance(self.timer, _NoTimer) else self.timer
        # trunk batches are independent: with n_streams > 1 they alternate between CUDA streams, so that the HBM-bound
        # kernels of one batch (max-pools, MAC) can overlap the tensor-bound convolutions of another
        main = torch.cuda.current_stream()
        side = self._side_streams(dev) if self.n_streams > 1 else None
        if side:
            ready = torch.cuda.Event()
            ready.record(main)
        for j, i in enumerate(range(0, K, self.trunk_batch)):
            st = side[j % len(side)] if side else main
            if ensure_crops is not None:
                ensure_crops(min(K, i + self.trunk_batch))
                if side:
                    ready = torch.cuda.Event()
                    ready.record(main)
            if side:
                st.wait_event(ready)
            with torch.cuda.stream(st):
                f4, f5 = self.encoder.features(crops[i:i + self.trunk_batch], normalized=True, pooled_f4=not self.rmac_levels)
                if self.rmac_levels and regions is None:
                    regions = [ops.rmac_regions(f.shape[2], f.shape[3], self.rmac_levels) for f in (f4, f5)]
                esz = f4.element_size()
                nb = float(f4.numel() + f5.numel()) * esz + 4.0 * D * f4.shape[0]
                with self.timer.span("region_pool_embed", nb):
                    ops.region_pool_embed([f4, f5], regions=regions, per_region_l2=bool(self.rmac_levels), out=emb[i:i + self.trunk_batch])
            self.launches += 1
        if ensure_crops is not None:
            ensure_crops(K + 1)  # groups without crops: nothing to launch, but the staging buffers are released
        if side:
            for st in side:
                crops.record_stream(st)
                emb.record_stream(st)
                main.wait_stream(st)
        self.launches += getattr(self.encoder, "own_launches", 0) - own0  # the library's max-pool launches inside the trunk
        G = self.gallery.G
        k = min(self.k, G)
        with self.timer.span("rows_prepare", 6.0 * K * D):
            qb, qinv = ops.rows_prepare(emb, normalize=True)
        self.launches += 1
        with self.timer.span("sim_topk", 2.0 * (K + G) * D + 8.0 * K * k, 2.0 * K * G * D):
            sim, idx = ops.sim_topk(qb, self.gallery.bf16, k, 0, emb if self.exact else None,
                                    self.gallery.f32 if self.exact else None, qinv if self.exact else None,
                                    self.gallery.inv_norm if self.exact else None)
        self.launches += 2 if self.exact else 1
        return {"boxes": boxes, "image_idx": image_idx, "counts": counts_h, "topk_idx": idx, "topk_sim": sim, "embedding": emb}
