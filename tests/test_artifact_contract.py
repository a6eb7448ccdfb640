"""The artifact side of the path: what ``rest_api`` does with the
pickle this path writes. ``recommend_tracks_for_track`` below restates
reference rest_api/app/main.py:224-254 (the service module itself imports fastapi_utils, which
this image lacks); the pickle is produced by the GPU path and by the CPU oracle's restatement
of the reference fold, and the service logic must answer identically on both."""
import os
import pickle
from collections import defaultdict

import numpy as np
import pytest
import torch

import kmls_b200
from kmls_b200 import api, miner, synth
from oracle import mlxtend_restated as M

pytestmark = pytest.mark.gpu
K_BEST_TRACKS = 10          # rest_api/app/main.py default


def recommend_tracks_for_track(songs_to_song_sets, seed_tracks):
    """rest_api/app/main.py:235-254 (the branch taken when a seed is known)."""
    songs_present = [t for t in seed_tracks if t in songs_to_song_sets]
    if not songs_present:
        return None                                   # the service falls back to its static list
    merged = defaultdict(int)
    for song in songs_present:
        for rec, value in songs_to_song_sets[song].items():
            merged[rec] = max(merged[rec], value)
    ranked = sorted(merged.items(), key=lambda x: x[1], reverse=True)     # stable: ties keep insertion order
    return [name for name, _ in ranked[:K_BEST_TRACKS]]


def _standin():
    d = synth.mpd_standin(seed=1)
    return synth.group_by_playlist(d["pid"], d["track_name"]), len(set(d["track_uri"]))


def test_pickle_round_trip_and_service_answers(tmp_path):
    playlists, total = _standin()
    recs, _, _ = kmls_b200.calculate_and_save_fp_growth_fast(playlists, 0.05, total_songs=total)
    want, _ = M.calculate_and_save_fp_growth_fast(playlists, 0.05)
    assert recs == want                                                  # dict ==, float64 values bit-equal
    # what save_pickle (main.py:136-145) writes and read_pickle_dict (rest_api/app/main.py:52-80) loads
    path = tmp_path / "recommendations.pickle"
    with open(path, "wb") as f:
        pickle.dump(recs, f)
    with open(path, "rb") as f:
        loaded = pickle.load(f)
    assert loaded == want
    assert all(type(k) is str and type(v) is dict for k, v in loaded.items())
    assert all(type(b) is str and type(x) is float for v in loaded.values() for b, x in v.items())
    rng = np.random.default_rng(0)
    keys = sorted(loaded)
    seeds = [[keys[int(i)] for i in rng.integers(0, len(keys), int(rng.integers(1, 4)))] for _ in range(200)]
    seeds += [["no such track"], ["no such track", keys[0]]]
    for seed in seeds:
        got, ref = recommend_tracks_for_track(loaded, seed), recommend_tracks_for_track(want, seed)
        if ref is None:
            assert got is None
            continue
        # same values in the same rank order; names may differ only inside a tie (the reference's own
        # tie order depends on Python's string hashing)
        merged = defaultdict(int)
        for s_ in seed:
            for r_, v_ in want.get(s_, {}).items():
                merged[r_] = max(merged[r_], v_)
        assert [merged[n_] for n_ in got] == [merged[n_] for n_ in ref]
        assert len(got) == len(ref)


def test_inner_dicts_are_value_descending_and_deterministic():
    playlists, total = _standin()
    a, _, _ = kmls_b200.calculate_and_save_fp_growth_fast(playlists, 0.05, total_songs=total)
    b, _, _ = kmls_b200.calculate_and_save_fp_growth_fast(dict(reversed(list(playlists.items()))), 0.05,
                                                          total_songs=total)
    assert a == b
    some = 0
    for key, inner in a.items():
        vals = list(inner.values())
        assert vals == sorted(vals, reverse=True)                        # support descending
        assert list(inner) == list(b[key])                               # same order whatever the input order
        # ties broken by name (= by track id, ids being ranks of the sorted names)
        for (n1, v1), (n2, v2) in zip(list(inner.items()), list(inner.items())[1:]):
            if v1 == v2:
                assert n1 < n2
                some += 1
    assert some > 0
    # with value-descending inner dicts the service's stable top-K is the same list on every run
    seed = [next(iter(a))]
    assert recommend_tracks_for_track(a, seed) == recommend_tracks_for_track(b, seed)


def test_save_rules_round_trip_matches_oracle(tmp_path):
    playlists, total = _standin()
    kmls_b200.calculate_and_save_fp_growth_fast(playlists, 0.05, total_songs=total, min_confidence=0.3)
    path = str(tmp_path / "rules.pickle")
    api.save_rules(path)
    with open(path, "rb") as f:
        data = pickle.load(f)
    assert set(data) == {"antecedents", "consequents", "support", "confidence", "lift"}
    n = len(data["support"])
    assert n > 0 and all(len(v) == n for v in data.values())
    assert type(data["antecedents"][0]) is tuple and type(data["antecedents"][0][0]) is str
    assert type(data["support"][0]) is float
    # against the restated mlxtend association_rules on the restated fpgrowth's itemsets
    _, itemsets = M.calculate_and_save_fp_growth_fast(playlists, 0.05)
    want = {(frozenset(a_), frozenset(c_)): (s_, cf, lf)
            for a_, c_, s_, cf, lf in M.association_rules(itemsets, 0.3)}
    got = {(frozenset(a_), frozenset(c_)): (s_, cf, lf)
           for a_, c_, s_, cf, lf in zip(data["antecedents"], data["consequents"], data["support"],
                                         data["confidence"], data["lift"])}
    assert got == want


def test_rules_to_host_is_vectorised_and_bounded(monkeypatch):
    off, items = synth.zipf_playlists(20000, 2000, mean_draws=35, alpha=1.0, seed=21)
    res, rules = kmls_b200.mine_ids(off, items, 2000, 0.008, 0.3)
    cols = api.rule_columns(res, rules)               # numpy columns, no per-rule Python
    assert len(cols["support"]) == rules.n_rules > 1000
    slow = miner.rules_to_host(res, rules)
    for r in range(0, rules.n_rules, max(1, rules.n_rules // 500)):
        lo, hi = cols["ante_off"][r], cols["ante_off"][r + 1]
        clo, chi = cols["cons_off"][r], cols["cons_off"][r + 1]
        a_, c_, s_, cf, lf = slow[r]
        assert frozenset(cols["ante_items"][lo:hi].tolist()) == a_
        assert frozenset(cols["cons_items"][clo:chi].tolist()) == c_
        assert (cols["support"][r], cols["confidence"][r], cols["lift"][r]) == (s_, cf, lf)
    monkeypatch.setattr(api, "RULES_HOST_LIMIT", 1000)
    with pytest.raises(kmls_b200._lib.KmlsError, match="host-memory bound"):
        api.rules_to_pinned(rules)


def test_best_tracks_percentile():
    playlists, _ = _standin()
    columns, offsets, ids = api.encode_transactions(list(playlists.values()))
    occ = np.bincount(ids, minlength=len(columns))
    tracks = api.most_frequent_tracks(columns, occ)
    kept = api.filter_best_tracks(tracks, 0.03)       # main.py:177-184 with TOP_TRACKS_SAVE_PERCENTILE = 0.03
    assert kept == tracks[:int(len(tracks) * 0.03)]
    assert len(kept) > 0 and kept[0]["count"] >= kept[-1]["count"]
