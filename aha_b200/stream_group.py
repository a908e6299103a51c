"""Several live streams on one GPU, stepped in lockstep ("per-GPU step batches its resident streams'
chunks into one M = 49 * S prefill").

The reference drives one stream per `LiveInferForBenchmark` object and one model call per frame
(test/inference.py:283-334); a stream's state never interacts with another's, so the frames of all resident streams
that are due at a step can share ONE `aha_prefill_chunk` call: the 13 GB of decoder weights are streamed once for the
whole group instead of once per stream.  Every stream keeps its own driver object (queues, cache, trigger state,
`debug_data_list`) exactly as in the reference; this class only changes who issues the model call.

    group = LiveInferGroup([LiveInferForBenchmark(model, start_ids, ...) for _ in range(8)])
    group.input_video_streams(list_of_uint8_videos)      # vision tower, frames of all streams interleaved per batch
    group.inference()                                    # fills every stream's debug_data_list
"""
from __future__ import annotations

from typing import List, Sequence

import torch

from .live_infer import LiveInferForBenchmark


class LiveInferGroup:
    def __init__(self, lives: Sequence[LiveInferForBenchmark], max_frames: int = 4096):
        if not lives:
            raise ValueError("LiveInferGroup needs at least one stream")
        self.lives: List[LiveInferForBenchmark] = list(lives)
        self.model = self.lives[0].model
        if any(lv.model is not self.model for lv in self.lives):
            raise ValueError("all streams of a group must share one model handle")
        self._scores = torch.zeros(max_frames, len(self.lives), 3, dtype=torch.float32, device=self.model.device)
        self._max_frames = max_frames
        self.reset_scores()

    def reset_scores(self):
        self._n_steps = 0
        self._rows = [[] for _ in self.lives]      # per stream: (step, column) of every frame it scored
        self._flushed = [0 for _ in self.lives]

    def reset(self):
        for lv in self.lives:
            lv.reset()
        self.reset_scores()

    # ---- vision ----
    def input_video_streams(self, videos: Sequence[torch.Tensor], batch_size: int = 32):
        """videos[s]: uint8 [T_s, 3, S, S] of stream s.  Frames are embedded `batch_size` per stream at a time with the
        streams interleaved, so that frame f of all streams is one contiguous [n_streams, q, H] block - the batched
        prefill call then takes it as is (no gather on the hot path)."""
        if len(videos) != len(self.lives):
            raise ValueError("one video per stream")
        dev, ns = self.model.device, len(self.lives)
        q, H = self.lives[0].frame_num_tokens, self.lives[0].hidden_size
        T = [int(v.shape[0]) for v in videos]
        for b in range(0, max(T), batch_size):
            n = [max(0, min(batch_size, t - b)) for t in T]
            full = min(n)
            if full > 0:
                # [full, ns, 3, S, S]: frame-major, stream-minor
                frames = torch.stack([v[b:b + full].to(dev, non_blocking=True) for v in videos], dim=1)
                emb = self.model.visual_embed(frames.reshape(-1, *frames.shape[2:])).view(full, ns, q, H)
                for f in range(full):
                    for s, lv in enumerate(self.lives):
                        lv.frame_embeds_queue.append(((b + f) / lv.frame_fps, emb[f, s]))
            for s, lv in enumerate(self.lives):       # ragged tails: the streams that are longer than the shortest one
                if n[s] > full:
                    e = self.model.visual_embed(videos[s][b + full:b + n[s]].to(dev, non_blocking=True)).view(-1, q, H)
                    for f in range(e.shape[0]):
                        lv.frame_embeds_queue.append(((b + full + f) / lv.frame_fps, e[f]))

    # ---- the loop ----
    def step(self) -> int:
        """One frame of every stream that still has one queued; returns how many streams stepped."""
        active = [i for i, lv in enumerate(self.lives) if lv.frame_embeds_queue]
        if not active:
            return 0
        if self._n_steps >= self._max_frames:
            raise RuntimeError("group score buffer full: raise max_frames")
        for i in active:                                  # queries that are due go first, per stream (test/inference.py:294-295)
            lv = self.lives[i]
            if lv.query_queue and lv.video_time >= lv.query_queue[0][0]:
                lv._encode_query()
        chunks = [self.lives[i]._next_frame_chunk() for i in active]
        # streams whose chunks have the same length share one call (steady state: all of them, q = frame_num_tokens)
        by_len = {}
        for j, (_, x, _) in enumerate(chunks):
            by_len.setdefault(int(x.shape[1]), []).append(j)
        col = 0
        row = self._scores[self._n_steps]
        for qlen, members in by_len.items():
            xs = [chunks[j][1] for j in members]
            x = self._stack(xs)
            caches = [self.lives[active[j]].past_key_values for j in members]
            out = self.model(inputs_embeds=x, use_cache=True, past_key_values=caches, return_dict=True,
                             scores_out=row[col:col + len(members)])
            for k, j in enumerate(members):
                lv = self.lives[active[j]]
                lv._frame_done(chunks[j][0], caches[k])
                self._rows[active[j]].append((self._n_steps, col + k))
            col += len(members)
        self._n_steps += 1
        if any(self.lives[i]._wants_decisions for i in active):
            host = row.tolist()                           # one sync for the whole group, only when a decision is due
            for i in active:
                lv = self.lives[i]
                if lv._wants_decisions:
                    inf, rel, _ = host[self._rows[i][-1][1]]
                    lv._after_frame_scores(inf, rel)
        for i in active:
            self.lives[i].video_time += 1 / self.lives[i].frame_fps
        return len(active)

    @staticmethod
    def _stack(xs: List[torch.Tensor]) -> torch.Tensor:
        """[1, q, H] chunks -> [n, q, H]; free when they already are consecutive slices of one buffer."""
        if len(xs) == 1:
            return xs[0]
        x0 = xs[0]
        step = x0.shape[1] * x0.shape[2] * x0.element_size()
        if all(x.is_contiguous() and x.data_ptr() == x0.data_ptr() + k * step for k, x in enumerate(xs)):
            return x0.as_strided((len(xs), x0.shape[1], x0.shape[2]), (x0.shape[1] * x0.shape[2], x0.shape[2], 1))
        return torch.cat(xs, dim=0)

    def inference(self):
        while self.step():
            pass
        self.flush_scores()
        return [lv.debug_data_list for lv in self.lives]

    def flush_scores(self):
        """The single device->host read: every stream's (time, scores) records (test/inference.py:299)."""
        if self._n_steps == 0:
            return
        host = self._scores[:self._n_steps].cpu().tolist()
        for i, lv in enumerate(self.lives):
            rows = self._rows[i][self._flushed[i]:]
            n0 = len(lv.debug_data_list)
            for (st, c), t in zip(rows, lv._times[n0:n0 + len(rows)]):
                inf, rel, unc = host[st][c]
                lv.debug_data_list.append(dict(time=t, informative_score=inf, relevance_score=rel, uncertainty_score=unc))
            self._flushed[i] = len(self._rows[i])
