"""Chunk planner: which frames go into which T-frame sampling run.

The caller of the sampling path (``run_one_scene``, seva/eval.py:1324-1974) cuts a scene into runs of T frames.
This module restates that integer bookkeeping -- bit-exact against the reference on the committed fixtures
(tests/golden/planner.json, generated from the real reference by the committed golden-vector script):

* ``pad_indices`` / ``assemble``                      eval.py:44-96
* ``infer_prior_stats`` / ``infer_prior_inds``        eval.py:344-453
* ``compute_relative_inds``                           eval.py:456-490
* ``find_nearest_source_inds``                        eval.py:493-501
* ``chunk_input_and_test``                            eval.py:504-852

Design: a chunk is a list of ``Slot`` integers pairs (kind, index) rather than the reference's "!003" / ">017" /
"NULL" strings; the string form is produced only by ``labels`` for printing and for the parity tests.  The three
strategy families are separate small generators (``_plan_gt``, ``_plan_nearest``, ``_plan_interp``).  ``ChunkPlan``
is what the multi-GPU scheduler (scene.py) consumes: per chunk the source / target indices, their slots in the
T-frame window, and the padded selection maps.
"""
from __future__ import annotations

import collections
import math
import re
from dataclasses import dataclass, field

import numpy as np
import torch

from .geometry import get_camera_dist

INPUT, TEST, NULL = 0, 1, 2   # slot kinds: conditioning frame ("!"), frame to generate (">"), padding
Slot = tuple  # (kind, index)


def labels(chunk: list) -> list:
    """The reference's printable form of a chunk (eval.py:517,608: "!%03d", ">%03d", "NULL")."""
    return ["NULL" if k == NULL else f"{'!' if k == INPUT else '>'}{i:03d}" for k, i in chunk]


def _fill(chunk: list, T: int) -> list:
    return chunk + [(NULL, 0)] * (T - len(chunk))


# ----------------------------------------------------------------------------------------------------
# window padding (eval.py:44-96)
# ----------------------------------------------------------------------------------------------------
def pad_indices(input_indices: list, test_indices: list, T: int, padding_mode: str = "last"):
    """Fill the unused slots of a T-frame window by repeating the LAST input or test selection -- whichever kind
    owns the highest used slot (eval.py:59-72).  Returns (input slots, test slots, input_maps, test_maps); the maps
    are int arrays over the window, -1 where the slot belongs to the other kind."""
    assert padding_mode in ["last", "none"], "`first` padding is not supported yet."
    used = set(input_indices) | set(test_indices)
    free = [i for i in range(T) if i not in used] if padding_mode == "last" else []
    in_sel = list(range(len(input_indices)))
    te_sel = list(range(len(test_indices)))

    def extend(slots, sels):
        sels = sels + [sels[-1]] * len(free)
        slots = list(slots) + free
        order = np.argsort(slots)
        return [slots[j] for j in order], [sels[j] for j in order]

    if max(input_indices) > max(test_indices):
        input_indices, in_sel = extend(input_indices, in_sel)
    else:
        test_indices, te_sel = extend(test_indices, te_sel)
    n = T if padding_mode == "last" else len(input_indices) + len(test_indices)
    input_maps = np.full(n, -1)
    test_maps = np.full(n, -1)
    input_maps[input_indices] = in_sel
    test_maps[test_indices] = te_sel
    return input_indices, test_indices, input_maps, test_maps


def assemble(input, test, input_maps, test_maps):
    """Gather the window's rows from the per-kind tensors using the ``pad_indices`` maps (eval.py:85-96)."""
    T = len(input_maps)
    im, tm = input_maps != -1, test_maps != -1
    assert np.logical_xor(im, tm).all()
    out = torch.zeros_like(test[-1:]).repeat_interleave(T, dim=0)
    out[im] = input[input_maps[im]]
    out[tm] = test[test_maps[tm]]
    return out


# ----------------------------------------------------------------------------------------------------
# two-pass sampling: how many anchor ("prior") frames the first pass generates, and which (eval.py:344-453)
# ----------------------------------------------------------------------------------------------------
def infer_prior_stats(T, num_input_frames: int, num_total_frames: int, version_dict: dict) -> int:
    """Number of anchor frames.  May widen the context windows: ``version_dict["T"]`` is rewritten in place when the
    input is semi-dense (>= ``num_input_semi_dense`` views), as the reference does (eval.py:387,419)."""
    opt = version_dict["options"]
    strategy = opt.get("chunk_strategy", "nearest")
    T1, T2 = (T[0], T[1]) if isinstance(T, (list, tuple)) else (T, T)
    semi_dense = num_input_frames >= opt.get("num_input_semi_dense", 9)
    ratio = opt.get("num_prior_frames_ratio", 1.0)
    floor_n = opt.get("num_prior_frames", 0)
    if not strategy.startswith("interp"):
        n = max(T1 - num_input_frames, floor_n)
        if semi_dense:
            version_dict["T"] = [n + num_input_frames, T2]
        return n
    # interp*: every second-pass window holds two anchors, so X + 1 anchors bound X windows of targets
    per_window = T2 - 2
    if not semi_dense and "gt" in strategy:
        per_window -= num_input_frames
    n = math.ceil(num_total_frames / per_window * ratio) + 1
    if n + num_input_frames < T1:
        n = T1 - num_input_frames
    n = max(n, floor_n)
    if semi_dense:
        version_dict["T"] = [n + num_input_frames, T2 + num_input_frames if "gt" in strategy else T2]
    return n


def infer_prior_inds(c2ws, num_prior_frames: int, input_frame_indices, options: dict) -> np.ndarray:
    """Anchor positions along the trajectory: evenly spread over the non-input frames (``interp*``) or greedy
    farthest-in-index-space from everything chosen so far (otherwise)."""
    n_frames = c2ws.shape[0]
    if options.get("chunk_strategy", "nearest").startswith("interp"):
        free = np.array([i for i in range(n_frames) if i not in input_frame_indices])
        pick = np.ceil(np.linspace(0, free.shape[0] - 1, num_prior_frames, endpoint=True)).astype(int)
        return np.sort(free[pick])
    chosen: list = []
    while len(chosen) < num_prior_frames:
        taken = np.concatenate([np.array(input_frame_indices), np.array(chosen)])
        gap = np.abs(np.arange(n_frames)[None] - taken[:, None]).min(0)
        chosen.append(np.argsort(gap)[-1])
    return np.sort(chosen)


def compute_relative_inds(source_inds, target_inds) -> list:
    """Fractional position of every target index on the axis spanned by ``source_inds`` (linear inter- /
    extrapolation in index space); used to place trajectory frames between anchors."""
    assert len(source_inds) > 2
    src = np.asarray(source_inds)
    n = len(src)
    out = []
    for ind in target_inds:
        hit = np.where(src == ind)[0]
        if len(hit):
            out.append(int(hit[0]))
        elif ind < src[0]:
            out.append(-((src[0] - ind) / (src[1] - src[0])))
        elif ind > src[-1]:
            out.append(n + ((ind - src[-1]) / (src[-1] - src[-2])))
        else:
            lo = int(np.where(src < ind)[0][-1])
            hi = int(np.where(src > ind)[0][0])
            out.append(lo + (ind - src[lo]) / (src[hi] - src[lo]) * (hi - lo))
    return out


def find_nearest_source_inds(source_c2ws, target_c2ws, nearest_num: int = 1, mode: str = "translation") -> np.ndarray:
    """[n_target, nearest_num] indices of the closest source cameras (eval.py:493-501)."""
    d = get_camera_dist(source_c2ws, target_c2ws, mode=mode).cpu().numpy()
    return np.argsort(d, axis=0).T[:, :nearest_num]


# ----------------------------------------------------------------------------------------------------
# chunking strategies (eval.py:504-785)
# ----------------------------------------------------------------------------------------------------
def _vote(rows: np.ndarray, k: int, exclude=None) -> list:
    """The k most frequent entries of ``rows`` (ties in first-seen order, ``Counter.most_common``)."""
    items = [v for v in rows.flatten().tolist() if exclude is None or v != exclude]
    return [v for v, _ in collections.Counter(items).most_common(k)]


def _plan_gt(T: int, n_gt: int, N: int, strategy: str, options: dict, test_c2ws) -> list:
    """``gt`` / ``gt-ltr`` / ``gt-nearest``: every window starts with all the real input views; after the first
    window ``gt-*`` also conditions on frames generated by EARLIER windows ("pseudo" inputs, numbered after the
    real ones), which is what chains first-pass windows together."""
    room = T - n_gt
    chunks, seen = [], 0
    while seen < N:
        chunk = [(INPUT, i) for i in range(n_gt)]
        n_pseudo = 0
        if strategy != "gt" and seen > 0:
            ratio = options.get("pseudo_num_ratio", 0.33)
            left = N - seen
            n_pseudo = math.ceil(room * ratio) if left >= math.floor(room * ratio) else room - left
            n_pseudo = min(n_pseudo, options.get("pseudo_num_max", 10000))
            if "ltr" in strategy:      # the most recently generated frames
                picked = list(range(seen - n_pseudo, seen))
            elif "nearest" in strategy:  # generated frames nearest (rotation and translation) to the coming targets
                near = np.concatenate(
                    [find_nearest_source_inds(test_c2ws[:seen], test_c2ws[seen:], 1, mode) for mode in
                     ("rotation", "translation")], axis=1)
                want = n_pseudo
                while True:  # fewer distinct neighbours than wanted frees target slots, which changes the vote
                    voted = _vote(near[: room - want], n_pseudo - 1, exclude=seen - 1)
                    picked = np.concatenate([np.sort(voted).astype(int), [seen - 1]])  # the last frame always stays
                    if len(picked) >= want:
                        break
                    want = len(picked)
                n_pseudo = len(picked)
                picked = [int(v) for v in picked]
            else:
                raise NotImplementedError(f"Chunking strategy {strategy} for the first pass is not implemented.")
            chunk += [(INPUT, i + n_gt) for i in picked]
        new = range(seen, min(seen + room - n_pseudo, N))
        chunk += [(TEST, i) for i in new]
        seen += len(new)
        chunks.append(_fill(chunk, T))
    return chunks


def _plan_nearest_k(T: int, k: int, input_c2ws, test_c2ws) -> list:
    """``nearest-K``: windows of T-K consecutive targets, each conditioned on the K source views most often nearest
    (translation) to them."""
    assert k < T, f"Nearest number of {k} should be less than {T}."
    N = test_c2ws.shape[0]
    near = find_nearest_source_inds(input_c2ws, test_c2ws, nearest_num=k, mode="translation")
    chunks = []
    for i in range(0, N, T - k):
        src = np.sort(_vote(near[i:i + T - k], k))
        chunk = [(INPUT, int(s)) for s in src] + [(TEST, j) for j in range(i, min(i + T - k, N))]
        chunks.append(_fill(chunk, T))
    return chunks


def _plan_nearest(T: int, gt: list, input_c2ws, test_c2ws) -> list:
    """``nearest`` / ``nearest-gt``: targets are grouped by their nearest source view (translation); groups are
    packed greedily into windows in source order, each group led by its source view (unless that view is one of
    the always-present ``gt`` views), and split when a window fills up."""
    N = test_c2ws.shape[0]
    near = find_nearest_source_inds(input_c2ws, test_c2ws, nearest_num=1, mode="translation")[:, 0]
    groups: dict = {}
    for t, s in enumerate(near.tolist()):
        groups.setdefault(s, []).append(t)
    queue = sorted(groups)
    head = [(INPUT, i) for i in gt]
    chunks, chunk, seen = [], list(head), 0
    while seen < N:
        s = queue[0]
        lead = [] if s in gt else [(INPUT, s)]
        if len(chunk) == T - len(lead):  # no room for a lead view plus one target: close the window
            if chunk:
                chunks.append(_fill(chunk, T))
                chunk = list(head)
            continue
        cand = lead + [(TEST, t) for t in groups[s]]
        room = T - len(chunk)
        if len(cand) <= room:
            chunk += cand
            seen += len(groups[s])
            queue.pop(0)
        else:
            chunk += cand[:room]
            took = room - len(lead)
            seen += took
            groups[s] = groups[s][took:]
        if len(chunk) == T:
            chunks.append(chunk)
            chunk = list(head)
    if chunk and chunk != head:
        chunks.append(_fill(chunk, T))
    return chunks


def _plan_interp(T: int, gt: list, M: int, input_ords: list, test_ords: list, with_gt: bool, base: int) -> list:
    """``interp`` / ``interp-gt``: source views are ordered stops along the trajectory; a window walks from stop to
    stop, taking all targets between two stops (a stop's targets are never split), and the next window re-opens at
    the stop where the previous one closed.  ``base`` offsets the stop numbering past the real inputs."""
    N = len(test_ords)
    stops = np.array([0] + list(input_ords), dtype=float)  # a virtual stop 0 catches targets before the first view
    stops[-1] += 0.01                                       # keep a target that coincides with the last stop
    nxt = np.concatenate([stops[1:], [np.inf]])
    t = np.array(test_ords)[None]
    between = np.logical_and(stops[:, None] <= t, nxt[:, None] > t)  # [M + 1, N]
    assert (between.sum(1) <= T - 2).all(), (
        "More anchor frames need to be sampled during the first pass to ensure "
        f"#target frames during each forward in the second pass will not exceed {T - 2}.")
    if stops[1] <= t[0, 0]:
        assert not between[0].any()
    if stops[-1] >= t[0, -1]:
        assert not between[-1].any()
    head = [(INPUT, i) for i in gt] if with_gt else []
    chunk = list(head) + [(TEST, j) for j in np.flatnonzero(between[0]).tolist()]
    between = between[1:]
    chunks, i = [], 0
    chunk.append((INPUT, base))
    while i < len(between):
        members = np.flatnonzero(between[i]).tolist()
        if not members:
            i += 1
            continue
        more_stops = i + 1 < M
        if len(members) + more_stops <= T - len(chunk):
            chunk += [(TEST, j) for j in members]
            i += 1
            if more_stops:
                chunk.append((INPUT, i + base))
        else:
            chunks.append(_fill(chunk, T))
            chunk = list(head) + [(INPUT, i + base)]
    if len(chunk) > 1:
        chunks.append(_fill(chunk, T))
    return chunks


def plan_chunks(T: int, input_c2ws, test_c2ws, input_ords, test_ords, options: dict, task: str = "img2img",
                chunk_strategy: str = "gt", gt_input_inds: list = ()) -> list:
    """Slot-list form of ``chunk_input_and_test``: a list of windows, each a list of T (kind, index) slots."""
    gt = list(gt_input_inds)
    M, N = input_c2ws.shape[0], test_c2ws.shape[0]
    if chunk_strategy.startswith("gt"):
        assert len(gt) < T, (f"Number of gt input frames {len(gt)} should be less than {T} when `gt` chunking "
                             "strategy is used.")
        assert list(range(M)) == gt, "All input_c2ws should be gt when `gt` chunking strategy is used."
        return _plan_gt(T, len(gt), N, chunk_strategy, options, test_c2ws)
    if chunk_strategy.startswith("nearest"):
        m = re.match(r"^nearest-(\d+)$", chunk_strategy)
        if m:
            return _plan_nearest_k(T, int(m.group(1)), input_c2ws, test_c2ws)
        return _plan_nearest(T, gt if "gt" in chunk_strategy else [], input_c2ws, test_c2ws)
    if chunk_strategy.startswith("interp"):
        assert input_ords is not None and test_ords is not None, (
            "When using `interp` chunking strategy, ordering of input and test frames should be provided.")
        base = 0
        input_ords = list(input_ords)
        if "img2trajvid" in task:  # the real inputs have no position on the trajectory: anchors only
            assert list(range(len(gt))) == gt, "`img2trajvid` task should put `gt_input_inds` in start."
            input_ords = [input_ords[i] for i in range(M) if i not in gt]
            M, base = len(input_ords), len(gt)
        return _plan_interp(T, gt, M, input_ords, list(test_ords), "gt" in chunk_strategy, base)
    raise NotImplementedError


def split_chunk(chunk: list):
    """(input indices, input slots, test indices, test slots) of one window.  A slot is the FIRST position holding
    an equal entry (``list.index``, eval.py:795-800): ``interp-gt`` windows outside img2trajvid can list a view
    twice, and both occurrences then report the first position."""
    first = {}
    for pos, s in enumerate(chunk):
        first.setdefault(s, pos)
    ins = [s for s in chunk if s[0] == INPUT]
    tes = [s for s in chunk if s[0] == TEST]
    return [i for _, i in ins], [first[s] for s in ins], [i for _, i in tes], [first[s] for s in tes]


def chunk_input_and_test(T, input_c2ws, test_c2ws, input_ords, test_ords, options, task: str = "img2img",
                         chunk_strategy: str = "gt", gt_input_inds: list = []):
    """Reference-shaped entry point (eval.py:504-852): (chunk labels, input indices, input slots, test indices,
    test slots), one list entry per window."""
    chunks = plan_chunks(T, input_c2ws, test_c2ws, input_ords, test_ords, options, task, chunk_strategy,
                         gt_input_inds)
    parts = [split_chunk(c) for c in chunks]
    if options.get("sampler_verbose", True):
        print("\nchunks:")
        for c in chunks:
            print(", ".join(labels(c)))
    return ([labels(c) for c in chunks], [p[0] for p in parts], [p[1] for p in parts], [p[2] for p in parts],
            [p[3] for p in parts])


# ----------------------------------------------------------------------------------------------------
# what the scheduler consumes
# ----------------------------------------------------------------------------------------------------
@dataclass
class Window:
    """One T-frame sampling run: which source rows / target rows it reads and where they sit in the window."""
    input_inds: list          # rows of the source table (real inputs, then frames generated so far)
    test_inds: list           # rows of the target table this run generates
    test_sels: list           # window slots of the generated frames (before padding)
    input_slots: list         # window slots conditioned on, after ``pad_indices``
    input_maps: np.ndarray    # [T] row of ``input_inds`` per slot, -1 elsewhere
    test_maps: np.ndarray     # [T] row of ``test_inds`` per slot, -1 elsewhere
    needs: list = field(default_factory=list)  # earlier windows whose output this one conditions on

    @property
    def mask(self) -> np.ndarray:
        m = np.zeros(len(self.input_maps), dtype=bool)
        m[self.input_slots] = True
        return m


def make_windows(T: int, input_inds_per_chunk, input_sels_per_chunk, test_inds_per_chunk, test_sels_per_chunk,
                 n_real_inputs: int | None = None, padding_mode: str = "last") -> list:
    """Windows with their dependency edges.  With ``n_real_inputs`` the source table is [real inputs; frames
    generated by earlier windows in generation order] (first pass / one pass, eval.py:1519-1527): source row r >=
    n_real_inputs is frame r - n_real_inputs of the concatenated earlier outputs.  ``None``: the source table is
    fixed (second pass) and windows are independent."""
    wins, produced = [], []  # produced[k] = window that generated the k-th frame
    for ii, isel, ti, tsel in zip(input_inds_per_chunk, input_sels_per_chunk, test_inds_per_chunk,
                                  test_sels_per_chunk):
        slots, _, imap, tmap = pad_indices(isel, tsel, T=T, padding_mode=padding_mode)
        needs = [] if n_real_inputs is None else sorted({produced[r - n_real_inputs] for r in ii if r >= n_real_inputs})
        wins.append(Window(list(ii), list(ti), list(tsel), list(slots), imap, tmap, needs))
        produced += [len(wins) - 1] * len(ti)
    return wins
