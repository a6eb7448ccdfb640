"""Deterministic synthetic transaction sets for the shapes BASELINE.json names.

The reference's transaction CSVs (``datasets/2023_spotify_ds1.csv``, ``..._ds2.csv``) are
absent from the checkout, so every benchmark and parity run uses these
stand-ins. Playlists are produced directly in the hand-off layout the device path takes:
CSR by playlist (``offsets[int64, N+1]``, ``items[int32, nnz]``), which is what
``list(playlist_dict.values())`` (machine-learning/main.py:265) is once names are ids.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np


_ZIPF_CHUNK = 1 << 16  # playlists per independently seeded chunk (part of the definition)


def _zipf_chunk(args):
    n_chunk, n_tracks, mean_draws, alpha, seed, stream = args
    rng = np.random.Generator(np.random.PCG64(seed).jumped(stream + 1))
    perm = np.random.Generator(np.random.PCG64(seed + 1)).permutation(n_tracks).astype(np.int64)
    w = np.arange(1, n_tracks + 1, dtype=np.float64) ** (-alpha)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    lens = np.clip(rng.poisson(mean_draws, n_chunk), 5, 250).astype(np.int64)
    u = rng.random(int(lens.sum()))
    ranks = np.minimum(np.searchsorted(cdf, u, side="left"), n_tracks - 1)
    pid = np.repeat(np.arange(n_chunk, dtype=np.int64), lens)
    key = np.unique(pid * n_tracks + perm[ranks])
    counts = np.bincount(key // n_tracks, minlength=n_chunk)
    return counts, (key % n_tracks).astype(np.int32)


def zipf_playlists(n_playlists: int, n_tracks: int, mean_draws: float = 72.0,
                   alpha: float = 1.0, seed: int = 0, shard: int = 0,
                   workers: int = 1) -> Tuple[np.ndarray, np.ndarray]:
    """Zipfian playlists.

    Per playlist: L = clip(Poisson(mean_draws), 5, 250) draws of a popularity rank by
    inverse CDF on weights r^-alpha; duplicates inside a playlist are dropped; rank r is
    mapped to a track id by a fixed permutation (seed+1) so ids are not popularity-sorted.
    Playlists are generated in chunks of 65,536, chunk c of shard s from the independent
    stream ``PCG64(seed).jumped(s * 65536 + c + 1)``: the result does not depend on
    ``workers`` (processes used to generate chunks in parallel), and each rank of a
    multi-GPU run can generate its own ``shard``.

    Returns (offsets int64[N+1], items int32[nnz]); items of a playlist are sorted by id.
    """
    jobs = []
    for c, p0 in enumerate(range(0, n_playlists, _ZIPF_CHUNK)):
        jobs.append((min(_ZIPF_CHUNK, n_playlists - p0), n_tracks, mean_draws, alpha, seed,
                     shard * 65536 + c))
    if workers > 1 and len(jobs) > 1:
        import concurrent.futures as cf
        with cf.ProcessPoolExecutor(max_workers=min(workers, len(jobs))) as ex:
            parts = list(ex.map(_zipf_chunk, jobs))
    else:
        parts = [_zipf_chunk(j) for j in jobs]
    offsets = np.zeros(n_playlists + 1, dtype=np.int64)
    if parts:
        np.cumsum(np.concatenate([c for c, _ in parts]), out=offsets[1:])
        items = np.concatenate([i for _, i in parts])
    else:
        items = np.zeros(0, dtype=np.int32)
    return offsets, items


def track_names(n_tracks: int) -> List[str]:
    """Names for synthetic ids; zero-padded so that sorted(names) keeps id order."""
    return [f"t{i:07d}" for i in range(n_tracks)]


def mpd_standin(seed: int = 1, n_playlists: int = 2246, n_uris: int = 2171,
                n_shared_names: int = 48, body_draws: float = 112.0, alpha: float = 0.5,
                n_hits: int = 16, hit_hi: float = 0.30, hit_lo: float = 0.10,
                vocabulary: List[str] | None = None) -> Dict[str, np.ndarray]:
    """Stand-in for ``2023_spotify_ds{1,2}.csv`` shaped after what the reference's report
    logs for ds2: 2,246 playlists, 2,171 track URIs of which 48 names are
    shared by two URIs, about 240k rows. Popularity: Zipf(alpha) body plus ``n_hits`` hit
    tracks each present in a playlist with probability hit_hi..hit_lo, calibrated so that the
    number of names at or above 3 / 5 / 10 % support (about 1.6k / 0.57k / 0.14k) follows the
    report's sweep plot (about 1.57k / 0.76k / 0.12k) and so that frequent pairs and triples
    exist at MIN_SUPPORT = 0.05 (independent draws alone give none).

    Returns the columns the hot path's callers use: ``pid`` (int64), ``track_uri`` (str),
    ``track_name`` (str), one row per (playlist, uri) occurrence, rows shuffled.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    n_names = n_uris - n_shared_names
    if vocabulary is None:
        vocabulary = [f"Track {i:04d}" for i in range(n_names)]
    else:
        seen = dict.fromkeys(vocabulary)
        vocabulary = list(seen)[:n_names]
        vocabulary += [f"Track {i:04d}" for i in range(n_names - len(vocabulary))]
    uri_name = np.arange(n_uris)
    uri_name[n_names:] = rng.choice(n_names, n_shared_names, replace=False)
    w = np.arange(1, n_uris + 1, dtype=np.float64) ** (-alpha)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    perm = rng.permutation(n_uris)
    lens = np.clip(rng.poisson(body_draws, n_playlists), 5, 250)
    u = rng.random(int(lens.sum()))
    uri = perm[np.minimum(np.searchsorted(cdf, u), n_uris - 1)]
    pid = np.repeat(np.arange(n_playlists, dtype=np.int64), lens)
    hit = rng.random((n_playlists, n_hits)) < np.linspace(hit_hi, hit_lo, n_hits)
    hp, hh = np.nonzero(hit)
    pid = np.concatenate([pid, hp])
    uri = np.concatenate([uri, perm[hh]])
    key = np.unique(pid * n_uris + uri)
    pid, uri = key // n_uris, key % n_uris
    order = rng.permutation(len(pid))
    pid, uri = pid[order], uri[order]
    names = np.array(vocabulary, dtype=object)[uri_name[uri]]
    uris = np.array([f"spotify:track:{i:022d}" for i in range(n_uris)], dtype=object)[uri]
    return {"pid": pid * 7 + 1000, "track_uri": uris, "track_name": names}


def group_by_playlist(pid: np.ndarray, names: np.ndarray) -> Dict[int, List[str]]:
    """What ``group_tracks_by_playlist_and_generate_homogeneous_data`` returns
    (machine-learning/main.py:195-207): ``{pid: [track_name, ...]}``, duplicates kept."""
    order = np.argsort(pid, kind="stable")
    pid_s, names_s = pid[order], names[order]
    cuts = np.flatnonzero(np.diff(pid_s)) + 1
    starts = np.concatenate([[0], cuts])
    ends = np.concatenate([cuts, [len(pid_s)]])
    return {int(pid_s[a]): list(names_s[a:b]) for a, b in zip(starts, ends)}
