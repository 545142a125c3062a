"""Host-side mirror of the reference's ``src/model.py`` API for the inference hot path:
``WrappedModel`` (:24-33), ``SpeakerEncoder`` (:36-135) and ``ModelHandling`` (:138-792) with the
same constructor arguments, method names, return conventions and checkpoint layout.

What changes is how the work is issued: the reference embeds one file per forward (batch =
num_eval crops) and scores one trial per Python iteration with a host sync each; here crops of
many files are packed into large batches for the libsvb encoder, H2D copies are double-buffered
on a side stream, and all trials are scored by batched kernels.  Training (``fit``), ONNX export
and plotting are out of scope and raise.
"""
import csv
import itertools
import os
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ecapa_tdnn, feature, raw3_ecapa, rawnet2_custom
from . import utils as svb_utils
from . import _native as nv
from .audio_loader import crop_plan, loadWAV, read_pcm

_ENCODERS = {"ECAPA_TDNN": ecapa_tdnn.MainModel, "RawNet2_custom": rawnet2_custom.MainModel,
             "Raw3_ECAPA": raw3_ecapa.MainModel}
# losses of the reference whose constructor sets test_normalize = False (src/losses/Prototypical.py:21)
# or leaves it unset (AttributeError -> False at src/model.py:90-94); every other loss sets True.
_NO_TEST_NORMALIZE = {"Prototypical", "AngularPenaltySMLoss", "CircleLoss", "CircleLossV2", "GE2ELoss",
                      "ProxyAnchor", "ProxyNCA", "adasoftmax_v1", "adasoftmax_v2"}


class WrappedModel(nn.Module):
    def __init__(self, model):
        super().__init__()
        self.module = model

    def forward(self, x, label=None):
        return self.module(x, label)


class _LossState(nn.Module):
    """Keeps the ``__L__.*`` tensors of a checkpoint so saveParameters round-trips them (training-only weights)."""

    def __init__(self, test_normalize):
        super().__init__()
        self.test_normalize = test_normalize

    def adopt(self, name, tensor):
        mod = self
        *path, leaf = name.split(".")
        for part in path:
            if not hasattr(mod, part):
                mod.add_module(part, nn.Module())
            mod = getattr(mod, part)
        mod.register_buffer(leaf, tensor.detach().clone())


class SpeakerEncoder(nn.Module):
    def __init__(self, model, criterion, classifier=None, optimizer=None, features="melspectrogram", device="cuda",
                 gpu=0, include_top=False, **kwargs):
        super().__init__()
        self.model, self.criterion, self.classifier, self.optimizer = model, criterion, classifier, optimizer
        self.gpu = gpu
        if str(device) != "cuda":
            raise RuntimeError("speakerverification_b200 runs on CUDA (sm_100a) only; device=%r has no code path" % (device,))
        self.device = torch.device(f"{device}:{self.gpu}")
        self.n_mels = kwargs["n_mels"]
        self.features = features.lower()
        if self.features == "melspectrogram":
            self.compute_features = feature.melspectrogram(**kwargs).to(self.device)
        elif self.features == "raw":
            self.compute_features = None                     # src/model.py:64-65: raw-waveform encoders
        else:
            raise NotImplementedError("features=%r: only 'melspectrogram' and 'raw' are on the accelerated path" % features)
        name = self.model["name"]
        if not isinstance(name, str) or name not in _ENCODERS:
            raise NotImplementedError("model %r is outside the accelerated hot path (available: %s)" % (name, sorted(_ENCODERS)))
        self.__S__ = _ENCODERS[name](nOut=self.model["nOut"], features=self.features, device=self.device,
                                     **kwargs).to(self.device)
        self.__L__ = _LossState(self.criterion["name"] not in _NO_TEST_NORMALIZE)
        self.test_normalize = self.__L__.test_normalize
        self.nPerSpeaker = kwargs["dataloader_options"]["nPerSpeaker"]
        self.include_top = include_top
        if include_top:
            raise NotImplementedError("include_top (classifier head) belongs to the training path")

    def engine(self):
        return self.__S__.engine(self.compute_features.frontend() if self.compute_features is not None else None)

    def forward(self, data, label=None):
        """[*, L] waveform crops (host or device) -> [B, nOut]; same reshape and ``squeeze`` as src/model.py:104-129."""
        if label is not None:
            raise NotImplementedError("the loss / training branch (src/model.py:130-135) is out of scope")
        data = data.reshape(-1, data.size()[-1]).to(self.device, dtype=torch.float32, non_blocking=True)
        return self.engine().forward(data).squeeze()


class ModelHandling(object):
    def __init__(self, encoder_model, optimizer="adam", callbacks="steplr", device="cuda", gpu=0, mixedprec=False,
                 **kwargs):
        self.kwargs = kwargs
        self.save_path = self.kwargs["save_folder"]
        self.audio_spec = self.kwargs["audio_spec"]
        self.__model__ = encoder_model
        self.model_name = self.__model__.module.model["name"]
        self.criterion = self.__model__.module.criterion["name"]
        self.gpu = gpu
        self.device = torch.device(f"{device}:{gpu}")
        self.mixedprec = mixedprec
        # rows (crops) per encoder launch and samples per ragged launch
        self.batch_rows = int(kwargs.get("svb_batch_rows", 592))       # 4 x 148 SMs (sweep: profiles/r01_batch_sweep.md)
        self.batch_samples = int(kwargs.get("svb_batch_samples", 296 * 48000))   # ragged launches: finer H2D double buffering
        self._copy_stream = None
        self._stage_bufs = None
        self._batch_bufs = None
        self._pool = None
        if self.gpu == 0:
            nb = sum(p.numel() for p in self.__model__.module.__S__.parameters())
            print("Model information:")
            print(f"- Initialized model: {self.model_name} - params: {nb:,} (libsvb, precision "
                  f"{self.__model__.module.__S__.precision})")
            print(f"- Embedding normalized:  {self.__model__.module.test_normalize}")

    # ------------------------------------------------------------------ out of scope
    def fit(self, *a, **k):
        raise NotImplementedError("training is out of scope for the B200 inference path")

    def export_onnx(self, *a, **k):
        raise NotImplementedError("ONNX export is out of scope")

    # ------------------------------------------------------------------ embedding pipeline
    def _engine(self):
        return self.__model__.module.engine()

    def embed_batches(self, waves, batch_rows=None, out=None):
        """Pipelined batch embedding: `waves` is a HOST fp32 tensor / array [R, L] (pinned for full overlap).
        H2D copies run on a side stream, double-buffered against the encoder; returns a CUDA [R, D] tensor."""
        if isinstance(waves, np.ndarray):
            waves = torch.from_numpy(waves)
        eng = self._engine()
        R, L = waves.shape
        bs = batch_rows or self.batch_rows
        if out is None:
            out = torch.empty((R, eng.n_out), dtype=torch.float32, device=self.device)
        if waves.is_cuda:
            for r0 in range(0, R, bs):
                eng.forward(waves[r0:r0 + bs], out=out[r0:r0 + bs])
            return out
        with torch.cuda.device(self.device):
            if self._copy_stream is None:
                self._copy_stream = torch.cuda.Stream(device=self.device)
            main = torch.cuda.current_stream(self.device)
            # the two device input buffers and their events live across calls (2 x 114 MB at 592 x 48 000): the
            # `freed` event of a buffer's last use in the PREVIOUS call is waited on like one of this call
            bb = self._batch_bufs
            if bb is None or bb["shape"][0] < bs or bb["shape"][1] != L:
                if bb is not None:
                    torch.cuda.synchronize(self.device)      # copies / kernels of the previous call may still use the old buffers
                self._batch_bufs = bb = None
                bb = self._batch_bufs = {"shape": (bs, L),
                                         "bufs": [torch.empty((bs, L), dtype=torch.float32, device=self.device)
                                                  for _ in range(2)],
                                         "ready": [torch.cuda.Event() for _ in range(2)],
                                         "freed": [None, None]}
            bufs, ready, freed = bb["bufs"], bb["ready"], bb["freed"]
            for i, r0 in enumerate(range(0, R, bs)):
                k = i & 1
                n = min(bs, R - r0)
                with torch.cuda.stream(self._copy_stream):
                    if freed[k] is not None:
                        self._copy_stream.wait_event(freed[k])
                    bufs[k][:n].copy_(waves[r0:r0 + n], non_blocking=True)
                    ready[k].record(self._copy_stream)
                main.wait_event(ready[k])
                eng.forward(bufs[k][:n], out=out[r0:r0 + n])
                if freed[k] is None:
                    freed[k] = torch.cuda.Event()
                freed[k].record(main)
        return out

    def _staging(self, t_dt, n_src, n_seg, n_dst):
        """Grow-only pinned / device staging buffers of the crop/stack pipeline, kept across calls (pinning is slow)."""
        st = self._stage_bufs
        need = (n_src, n_seg, n_dst)
        if st is None or st["dt"] != t_dt or any(x < y for x, y in zip(st["n"], need)):
            grow = need if st is None or st["dt"] != t_dt else tuple(max(x, y) for x, y in zip(st["n"], need))
            if st is not None:
                torch.cuda.synchronize(self.device)          # in-flight copies / kernels still use the old buffers
            self._stage_bufs = st = None
            st = {"dt": t_dt, "n": grow, "ready": [None, None], "freed": [None, None],
                  "h_pcm": [torch.empty(grow[0], dtype=t_dt).pin_memory() for _ in range(2)],
                  "h_seg": [torch.empty(grow[1], dtype=torch.uint8).pin_memory() for _ in range(2)],
                  "d_pcm": [torch.empty(grow[0], dtype=t_dt, device=self.device) for _ in range(2)],
                  "d_seg": [torch.empty(grow[1], dtype=torch.uint8, device=self.device) for _ in range(2)],
                  "d_wav": [torch.empty(grow[2], dtype=torch.float32, device=self.device) for _ in range(2)]}
            self._stage_bufs = st
        return st

    stage_trace = None   # set to a list to collect per-group host staging times of embed_signals

    def embed_signals(self, signals, num_eval):
        """GPU crop/stack pipeline (replaces loadWAV eval mode + test_data_loader stacking,
        src/processing/audio_loader.py:110-150, src/dataloader.py:275-292): `signals` is a list of 1-D int16 / float32
        HOST arrays.  Every utterance crosses PCIe once (int16 stays int16), is converted / wrap-padded on the device
        by svb_pcm_stage, and its num_eval crops are row offsets into that staged buffer -- no crop is copied.
        Launches are double-buffered: the H2D copy of group i+1 overlaps the encoder on group i.
        Returns CUDA [U, n, D] (n = num_eval, or 1 whole-utterance row when num_eval == 0)."""
        eng = self._engine()
        U = len(signals)
        n = num_eval if num_eval > 0 else 1
        out = torch.empty((U * n, eng.n_out), dtype=torch.float32, device=self.device)
        if U == 0:
            return out.view(0, n, eng.n_out)
        i16 = all(s.dtype == np.int16 for s in signals)
        if not i16:
            signals = [s.astype(np.float32) / 32768.0 if s.dtype == np.int16 else np.asarray(s, dtype=np.float32)
                       for s in signals]
        src_len = np.array([len(s) for s in signals], dtype=np.int64)
        dst_len, starts, L = crop_plan(src_len, self.audio_spec, num_eval)
        # groups of utterances per encoder launch
        # (plain Python ints and prefix sums: this prologue runs before the first byte moves, nothing overlaps it)
        dl = dst_len.tolist()
        groups, i = [], 0
        while i < U:
            j, rows, samples = i, 0, 0
            while j < U and (j == i or (rows + n <= self.batch_rows and
                                        (num_eval > 0 or samples + dl[j] <= self.batch_samples))):
                rows += n
                samples += dl[j]
                j += 1
            groups.append((i, j))
            i = j
        csrc = np.concatenate([[0], np.cumsum(src_len)])
        cdst = np.concatenate([[0], np.cumsum(dst_len)])
        max_src = max(int(csrc[b] - csrc[a]) for a, b in groups)
        max_dst = max(int(cdst[b] - cdst[a]) for a, b in groups)
        max_u = max(b - a for a, b in groups)
        t_dt = torch.int16 if i16 else torch.float32
        seg_dt = np.dtype([("src_off", "<i8"), ("dst_off", "<i8"), ("src_len", "<i4"), ("dst_len", "<i4")])
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        copy_stream = self._copy_stream
        main = torch.cuda.current_stream(self.device)
        st = self._staging(t_dt, max_src, max_u * seg_dt.itemsize, max_dst)
        h_pcm, h_seg, d_pcm, d_seg, d_wav = st["h_pcm"], st["h_seg"], st["d_pcm"], st["d_seg"], st["d_wav"]
        # `ready[k]` (H2D copies of slot k executed) and `freed[k]` (the encoder finished with slot k) persist with the
        # staging buffers: the first two groups of THIS call wait on the last uses of the PREVIOUS call, which may still
        # be queued (embed_signals returns without a synchronisation)
        ready, freed = st["ready"], st["freed"]
        if self._pool is None:
            # one stager + pinned-fill workers; the ranks of one box share its host cores
            local_world = int(os.environ.get("LOCAL_WORLD_SIZE", "1") or 1)
            self._pool = ThreadPoolExecutor(max(2, min(6, (os.cpu_count() or 8) // max(local_world, 1))))

        def fill(k, lo, hi, offs):
            hp = h_pcm[k].numpy()
            for u in range(lo, hi):
                hp[offs[u]:offs[u] + len(signals[u])] = signals[u]

        def stage(g):
            """Host side of group g (runs on a worker thread, one group ahead of the encoder): PCM of its utterances
            into pinned buffer g & 1 (parallel memcpy, numpy releases the GIL) and the H2D copies on the copy stream."""
            a, b = groups[g]
            k = g & 1
            t_in = time.perf_counter()
            if ready[k] is not None:
                ready[k].synchronize()                       # the pinned buffers' previous contents have been copied out
            t_sync = time.perf_counter()
            segs = np.zeros(b - a, dtype=seg_dt)
            segs["src_len"], segs["dst_len"] = src_len[a:b], dst_len[a:b]
            segs["src_off"] = np.concatenate([[0], np.cumsum(src_len[a:b])[:-1]])
            segs["dst_off"] = np.concatenate([[0], np.cumsum(dst_len[a:b])[:-1]])
            offs = {u: int(segs["src_off"][u - a]) for u in range(a, b)}
            cuts = np.linspace(a, b, num=min(5, b - a) + 1).astype(int)   # (10-way fill on 16 cores measured slower: 14.9 vs 16.0-16.5 k utt/s)
            list(self._pool.map(lambda lh: fill(k, lh[0], lh[1], offs), zip(cuts[:-1], cuts[1:])))
            nsrc = int(src_len[a:b].sum())
            t_fill = time.perf_counter()
            h_seg[k].numpy()[:segs.nbytes] = segs.view(np.uint8)
            with torch.cuda.device(self.device), torch.cuda.stream(copy_stream):
                if freed[k] is not None:
                    copy_stream.wait_event(freed[k])         # the previous user of d_pcm[k] / d_seg[k] has run
                d_pcm[k][:nsrc].copy_(h_pcm[k][:nsrc], non_blocking=True)
                d_seg[k][:segs.nbytes].copy_(h_seg[k][:segs.nbytes], non_blocking=True)
                if ready[k] is None:
                    ready[k] = torch.cuda.Event()
                ready[k].record(copy_stream)
            if self.stage_trace is not None:                 # diagnostics: (group, wait, fill, enqueue) in ms
                t_out = time.perf_counter()
                self.stage_trace.append((g, 1e3 * (t_sync - t_in), 1e3 * (t_fill - t_sync), 1e3 * (t_out - t_fill)))
            return segs

        fut = self._pool.submit(stage, 0)
        with torch.cuda.device(self.device):
            for g, (a, b) in enumerate(groups):
                k = g & 1
                segs = fut.result()
                if g + 1 < len(groups):
                    fut = self._pool.submit(stage, g + 1)    # freed[(g+1)&1] was recorded when group g-1 was enqueued
                main.wait_event(ready[k])
                nv.check(nv.lib().svb_pcm_stage(nv.ptr(d_pcm[k]), 1 if i16 else 0, nv.ptr(d_seg[k]), b - a,
                                                int(dst_len[a:b].max()), nv.ptr(d_wav[k]), nv.stream_ptr()))
                base = segs["dst_off"]
                if num_eval > 0:
                    offs = np.concatenate([base[u - a] + starts[u] for u in range(a, b)])
                    lens = np.full(len(offs), L, dtype=np.int64)
                else:
                    offs, lens = base, dst_len[a:b]
                eng.forward_rows(d_wav[k], offs, lens, out=out[a * n:b * n])
                if freed[k] is None:
                    freed[k] = torch.cuda.Event()
                freed[k].record(main)
        return out.view(U, n, eng.n_out)

    def _embed_ragged(self, signals):
        """Whole-utterance embeddings (num_eval == 0): every signal keeps its own length."""
        eng = self._engine()
        out = torch.empty((len(signals), eng.n_out), dtype=torch.float32, device=self.device)
        i = 0
        while i < len(signals):
            j, tot = i, 0
            while j < len(signals) and (j == i or tot + len(signals[j]) <= self.batch_samples):
                tot += len(signals[j])
                j += 1
            flat = torch.from_numpy(np.concatenate(signals[i:j]).astype(np.float32)).to(self.device, non_blocking=True)
            eng.forward_flat(flat, [len(s) for s in signals[i:j]], out=out[i:j])
            i = j
        return out

    def _load_crops(self, files, num_eval, num_workers=0):
        def one(f):
            return loadWAV(f, self.audio_spec, evalmode=True, augment=False, augment_options=[], num_eval=num_eval,
                           random_chunk=False)
        if num_workers and num_workers > 1:
            with ThreadPoolExecutor(num_workers) as ex:
                return list(ex.map(one, files))
        return [one(f) for f in files]

    def _embed_files(self, files, num_eval, num_workers=0):
        """list of paths -> CUDA [U, n, D] crop embeddings (n = num_eval, or 1 when num_eval == 0)."""
        if hasattr(self._engine(), "forward_rows"):
            if num_workers and num_workers > 1:
                with ThreadPoolExecutor(num_workers) as ex:
                    signals = list(ex.map(read_pcm, [str(Path(f)) for f in files]))
            else:
                signals = [read_pcm(str(Path(f))) for f in files]
            return self.embed_signals(signals, num_eval)
        crops = self._load_crops(files, num_eval, num_workers)      # fixed-length raw-waveform encoders (RawNet2)
        if num_eval == 0:
            return self._embed_ragged([c[0] for c in crops]).unsqueeze(1)
        allc = torch.from_numpy(np.concatenate(crops, axis=0))
        try:
            allc = allc.pin_memory()
        except RuntimeError:
            pass
        emb = self.embed_batches(allc)
        return emb.view(len(files), num_eval, -1)

    # ------------------------------------------------------------------ multi-GPU plumbing
    @staticmethod
    def shard(n_items, rank, world):
        """Round-robin utterance sharding (the DistributedSampler(shuffle=False) split of src/model.py:363-367)."""
        return list(range(rank, n_items, world))

    @staticmethod
    def gather_sharded(local, n_items, rank, world):
        """All-gather of per-rank [U_r, ...] tensors back into item order [n_items, ...] on every rank.
        Replaces all_gather_object of pickled CPU dicts (src/model.py:400-411) by one tensor collective."""
        import torch.distributed as dist
        per = (n_items + world - 1) // world
        pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
        parts = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(parts, pad)
        full = torch.empty((n_items,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        for r in range(world):
            idx = list(range(r, n_items, world))
            full[idx] = parts[r][:len(idx)]
        return full

    @staticmethod
    def shard_balanced(lengths, world):
        """Greedy length-balanced assignment for ragged utterances: longest first onto
        the least-loaded rank.  Returns `world` index arrays, each in ascending item order; identical on
        every rank (pure function of the lengths)."""
        lengths = np.asarray(lengths, dtype=np.int64)
        order = np.argsort(-lengths, kind="stable")
        loads = [0] * world
        parts = [[] for _ in range(world)]
        for i in order:
            r = min(range(world), key=lambda k: (loads[k], k))
            parts[r].append(int(i))
            loads[r] += int(lengths[i])
        return [np.asarray(sorted(p), dtype=np.int64) for p in parts]

    @staticmethod
    def gather_assigned(local, parts, rank, world):
        """All-gather for an explicit assignment `parts` (from shard_balanced): [len(parts[rank]), ...] per rank
        -> [n_items, ...] in item order on every rank."""
        import torch.distributed as dist
        n_items = int(sum(len(p) for p in parts))
        per = max(1, max(len(p) for p in parts))
        pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
        allp = torch.empty((world * per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(allp, pad)
        full = torch.empty((n_items,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        for r in range(world):
            if len(parts[r]):
                full[torch.as_tensor(parts[r], device=local.device)] = allp[r * per:r * per + len(parts[r])]
        return full

    # ------------------------------------------------------------------ evaluation from a trial list
    def evaluateFromList(self, listfilename, distributed, dataloader_options, cohorts_path="checkpoint/dump_cohorts.npy",
                         num_eval=10, scoring_mode="cosine", **kwargs):
        """Same contract as src/model.py:306-450: returns (scores, labels, trials); only rank 0 returns
        non-empty lists when distributed.  Trial file: ``label path1 path2`` per line."""
        rank, world = 0, 1
        if distributed:
            import torch.distributed as dist
            rank, world = dist.get_rank(), dist.get_world_size()
        self.__model__.eval()
        cohorts = None
        if cohorts_path is not None and scoring_mode == "norm":
            cohorts = np.load(cohorts_path)
        if cohorts_path is None:
            raise NotImplementedError("cohorts_path=None selects the training-validation distance branch (out of scope)")
        with open(listfilename) as f:
            lines = f.readlines()
        sep = "," if len(lines[0].split(",")) > 1 else " "
        start = 1 if sep == "," else 0
        files = list(itertools.chain(*[x.strip().split(sep)[-2:] for x in lines[start:]]))
        setfiles = sorted(set(files))
        print(">>>>Evaluation")
        mine = self.shard(len(setfiles), rank, world)
        workers = (dataloader_options or {}).get("num_workers", 0)
        emb = self._embed_files([setfiles[i] for i in mine], num_eval, workers)
        if distributed and world > 1:
            emb = self.gather_sharded(emb, len(setfiles), rank, world)
        if rank != 0:
            return [], [], []
        index = {str(Path(f)): i for i, f in enumerate(setfiles)}
        index.update({f: i for i, f in enumerate(setfiles)})
        labels, ia, ib, trials = [], [], [], []
        for line in lines:
            d = line.split()
            labels.append(int(d[0]))
            ia.append(index[d[1]])
            ib.append(index[d[2]])
            trials.append(d[1] + " " + d[2])
        scores = self._score(emb, ia, ib, scoring_mode, cohorts, top=kwargs.get("top", 200))
        return scores, labels, trials

    def _score(self, emb, ia, ib, scoring_mode, cohorts, top=200):
        ia = torch.tensor(ia, dtype=torch.int32, device=self.device)
        ib = torch.tensor(ib, dtype=torch.int32, device=self.device)
        s = svb_utils.score_trials(emb.contiguous(), ia, ib, scoring_mode, cohorts, top=top,
                                   normalize=self.__model__.module.test_normalize)
        return [float(v) for v in s.cpu().numpy()]

    # ------------------------------------------------------------------ testing from a CSV pair list
    def testFromList(self, test_list="evaluation_test.txt", thresh_score=0.5, distributed=False, dataloader_options=None,
                     cohorts_path=None, num_eval=10, scoring_mode="norm", output_file=None):
        """Same contract as src/model.py:455-554 (CSV with header in, result CSV out, 'a,b,score' strings back)."""
        self.__model__.eval()
        cohorts = None
        if cohorts_path is not None and scoring_mode == "norm":
            cohorts = np.load(cohorts_path)
        save_root = os.path.join(self.save_path, f"{self.model_name}/{self.criterion}/result")
        if output_file is None:
            output_file = test_list.replace(".txt", "_result.txt")
        write_file = Path(save_root, output_file) if os.path.split(output_file)[0] == "" else output_file
        print(">>>>TESTING...")
        print(f">>> Threshold: {thresh_score}")
        files, lines = [], []
        with open(Path(test_list), newline="") as rf:
            reader = csv.reader(rf, delimiter=",")
            next(reader, None)
            for row in reader:
                files += [row[0], row[1]]
                lines.append(row)
        setfiles = sorted(set(files))
        workers = (dataloader_options or {}).get("num_workers", 0)
        emb = self._embed_files(setfiles, num_eval, workers)
        index = {f: i for i, f in enumerate(setfiles)}
        scores = self._score(emb, [index[r[0]] for r in lines], [index[r[1]] for r in lines], scoring_mode, cohorts)
        results = []
        os.makedirs(os.path.dirname(str(write_file)) or ".", exist_ok=True)
        with open(write_file, "w", newline="") as wf:
            w = csv.writer(wf, delimiter=",")
            w.writerow(["audio_1", "audio_2", "pred_label", "score"])
            for row, score in zip(lines, scores):
                w.writerow([row[0], row[1], "1" if score >= thresh_score else "0", score])
                results.append(f"{Path(row[0]).name},{Path(row[1]).name},{score}")
        return results

    # ------------------------------------------------------------------ cohort / enrolment preparation
    def prepare(self, save_path=None, prepare_type="cohorts", num_eval=10, source=None, **kwargs):
        """src/model.py:559-670.  'cohorts': ``spkID path`` lines -> per-speaker mean of the L2-normalised crop
        embeddings of the first 3 files (not re-normalised), saved with np.save.  'embed' with a list of
        arrays -> mean embedding."""
        self.__model__.eval()
        if not source:
            raise ValueError("Please provide appropriate source!")
        if prepare_type == "cohorts":
            assert isinstance(source, str), "Please provide path to train metadata files"
            spk_files = {}
            with open(Path(source)) as f:
                for line in f.read().splitlines():
                    d = line.split()
                    if len(d) >= 2:
                        spk_files.setdefault(d[0], []).append(d[1])
            flat, owner = [], []
            for s, (spk, paths) in enumerate(spk_files.items()):
                for p in paths[:3]:
                    flat.append(p)
                    owner.append(s)
            emb = F.normalize(self._embed_files(flat, num_eval), p=2, dim=2)          # [files, n, D]
            owner_t = torch.tensor(owner, device=self.device)
            S, D = len(spk_files), emb.shape[-1]
            sums = torch.zeros((S, D), device=self.device).index_add_(0, owner_t, emb.sum(1))
            cnt = torch.zeros(S, device=self.device).index_add_(0, owner_t, torch.full((len(owner),), float(emb.shape[1]), device=self.device))
            cohort = (sums / cnt[:, None]).cpu().numpy()
            if save_path:
                np.save(save_path, cohort)
            return True
        if prepare_type == "embed" and isinstance(source, (str, Path)):
            # src/model.py:615-645: one directory per known speaker; bank [n, D, S] = per-speaker mean over its files
            # of the crop embeddings, speakers on the LAST axis (embeds.pt), classes = {index: directory name}.
            # All files go through the batched crop/stack pipeline in one pass (directories in sorted order).
            dirs = sorted(x for x in Path(source).iterdir() if x.is_dir())
            classes = {i: d.stem for i, d in enumerate(dirs)}
            flat, owner = [], []
            for i, d in enumerate(dirs):
                for f in sorted(d.glob("*.wav")):
                    flat.append(str(f))
                    owner.append(i)
            if not flat or len(set(owner)) != len(dirs):
                raise ValueError("every speaker directory under %s needs at least one .wav file" % source)
            emb = self._embed_files(flat, num_eval)                                    # [files, n, D]
            if self.__model__.module.test_normalize:
                emb = F.normalize(emb, p=2, dim=2)
            owner_t = torch.tensor(owner, device=self.device)
            S = len(dirs)
            sums = torch.zeros((S,) + tuple(emb.shape[1:]), device=self.device).index_add_(0, owner_t, emb)
            cnt = torch.bincount(owner_t, minlength=S).to(sums.dtype)
            embeds = (sums / cnt[:, None, None]).permute(1, 2, 0).contiguous().cpu()   # [n, D, S]
            if save_path:
                torch.save(embeds, Path(save_path, "embeds.pt"))
                np.save(str(Path(save_path, "classes.npy")), classes)
            return True
        if prepare_type == "embed" and isinstance(source, list):
            embs = [self.embed_utterance(a, num_eval=num_eval, normalize=self.__model__.module.test_normalize)
                    for a in source]
            mean_embed = torch.mean(torch.stack(embs), dim=0)
            if save_path:
                torch.save(mean_embed, Path(save_path, "embeds.pt"))
            return mean_embed
        raise NotImplementedError("prepare_type=%r with source type %s" % (prepare_type, type(source).__name__))

    def predict(self, sources, embeds, classes=None, num_eval=10, topk=3):
        """The ``--predict`` loop of src/inference.py:254-327 as one batched call: every file of ``sources`` is
        embedded (``embed_utterance`` semantics: ``num_eval`` crops, normalised when the loss asks for it), compared
        with the enrolment bank ``embeds`` [n, D, S] (tensor or path of embeds.pt) by the mean-over-crops L2 distance,
        and the ``topk`` nearest speakers are returned: dict(dist [Q, S], top_dist [Q, k], top_idx [Q, k],
        score = 1 - top_dist**2 / 2 (:293), names = classes[top_idx] when ``classes`` (dict or classes.npy path) is
        given).  The reference's console report (same_smallest_score / diff_biggest_score) is host logic on these."""
        from .utils import enrol_nearest
        if isinstance(embeds, (str, Path)):
            embeds = torch.load(embeds)
        if isinstance(classes, (str, Path)):
            classes = np.load(str(classes), allow_pickle=True).item()
        emb = self._embed_files([str(f) for f in sources], num_eval)                   # [Q, n, D] on the device
        if self.__model__.module.test_normalize:
            emb = F.normalize(emb, p=2, dim=2)
        dist, td, ti = enrol_nearest(emb.contiguous(), embeds.to(self.device), topk=topk)
        out = dict(dist=dist.cpu().numpy(), top_dist=td.cpu().numpy(), top_idx=ti.cpu().numpy())
        out["score"] = 1.0 - out["top_dist"] ** 2 / 2
        if classes is not None:
            out["names"] = [[classes[int(i)] for i in row] for row in out["top_idx"]]
        return out

    def embed_utterance(self, source, num_eval=20, normalize=False):
        """src/model.py:675-704: path or ndarray -> CPU tensor [num_eval, D]."""
        audio = loadWAV(source, self.audio_spec, evalmode=True, augment=False, augment_options=[], num_eval=num_eval,
                        random_chunk=False)
        inp = torch.from_numpy(np.ascontiguousarray(audio)).to(self.device)
        with torch.no_grad():
            embed = self._engine().forward(inp).detach().cpu()
        if normalize:
            embed = F.normalize(embed, p=2, dim=1)
        return embed

    # ------------------------------------------------------------------ checkpoints (reference layout)
    def saveParameters(self, path):
        torch.save(self.__model__.module.state_dict(), path)

    def loadParameters(self, path, show_error=True):
        """src/model.py:718-746: tolerant load (strips 'module.', skips unknown / mis-shaped keys).  Unknown
        ``__L__.*`` tensors are adopted so that saveParameters writes them back."""
        if not os.path.exists(path):
            raise FileNotFoundError("Model's path is not exists")
        self_state = self.__model__.module.state_dict()
        loaded = torch.load(path, map_location=self.device)
        for name, param in loaded.items():
            orig = name
            if name not in self_state:
                name = name.replace("module.", "")
                if name not in self_state:
                    if name.startswith("__L__.") and torch.is_tensor(param):
                        self.__model__.module.__L__.adopt(name[len("__L__."):], param)
                    elif show_error:
                        print("{} is not in the model.".format(orig))
                    continue
            if self_state[name].size() != param.size():
                if show_error:
                    print("Wrong parameter length: {}, model: {}, loaded: {}".format(orig, self_state[name].size(), param.size()))
                continue
            self_state[name].copy_(param)
