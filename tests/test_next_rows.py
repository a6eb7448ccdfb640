"""the callers either side of the hot path - ingest from the CSV columns,
the best-tracks list from the level-1 kernel, the support sweep with shared level-1/2 work."""
import numpy as np
import pandas as pd
import pytest
import torch

import kmls_b200
from kmls_b200 import api, miner, synth
from oracle import mlxtend_restated as M


def test_encode_frame_matches_the_dict_route():
    d = synth.mpd_standin(seed=3, n_playlists=300, n_uris=200, n_shared_names=7, body_draws=20, n_hits=4)
    playlists = synth.group_by_playlist(d["pid"], d["track_name"])
    cols_a, off_a, ids_a = api.encode_transactions(list(playlists.values()))
    cols_b, off_b, ids_b = api.encode_frame(d["pid"], d["track_name"])
    assert cols_a == cols_b and (off_a == off_b).all()
    # same multiset of ids inside every playlist (the dict route keeps file order too)
    assert (ids_a == ids_b).all()


def test_most_frequent_tracks_is_the_reference_group_count():
    d = synth.mpd_standin(seed=4, n_playlists=200, n_uris=150, n_shared_names=5, body_draws=15, n_hits=3)
    cols, off, ids = api.encode_frame(d["pid"], d["track_name"])
    occ = np.bincount(ids, minlength=len(cols))
    got = api.most_frequent_tracks(cols, occ)
    want = pd.Series(d["track_name"]).value_counts()
    assert {g["track_name"]: g["count"] for g in got} == want.to_dict()
    assert [g["count"] for g in got] == sorted((g["count"] for g in got), reverse=True)


@pytest.mark.gpu
def test_frame_route_and_histogram_on_gpu():
    d = synth.mpd_standin(seed=1)
    playlists = synth.group_by_playlist(d["pid"], d["track_name"])
    total = len(set(d["track_uri"]))
    a, _, (ma, _) = kmls_b200.calculate_and_save_fp_growth_fast(playlists, 0.05, total_songs=total)
    b, _, (mb, _) = kmls_b200.recommendations_from_frame(d["pid"], d["track_name"], 0.05, total_songs=total)
    assert a == b and ma == mb
    cols, off, ids = api.encode_frame(d["pid"], d["track_name"])
    dev = torch.device("cuda", 0)
    prep = miner.prepare(torch.from_numpy(off).to(dev), torch.from_numpy(ids).to(dev), len(cols), 0.05)
    got = api.most_frequent_tracks(cols, prep.occurrences)
    assert {g["track_name"]: g["count"] for g in got} == pd.Series(d["track_name"]).value_counts().to_dict()


@pytest.mark.gpu
def test_sweep_reuses_levels_1_and_2_and_matches_independent_runs():
    d = synth.mpd_standin(seed=2)
    playlists = synth.group_by_playlist(d["pid"], d["track_name"])
    total = len(set(d["track_uri"]))
    supports = [0.05, 0.06, 0.08, 0.1, 0.2]
    frame = kmls_b200.sweep_supports(playlists, supports, total_songs=total)
    assert list(frame["min_support"]) == supports
    for s, missing, n_sets in zip(frame["min_support"], frame["songs_without_recommendations"], frame["itemsets"]):
        recs, data = M.calculate_and_save_fp_growth_fast(playlists, s)
        assert missing == total - len(recs)
        assert n_sets == len(data)
    # itemset tables from the shared preparation equal the ones of a fresh run
    cols, off, ids = api.encode_transactions(list(playlists.values()))
    dev = torch.device("cuda", 0)
    d_off, d_ids = torch.from_numpy(off).to(dev), torch.from_numpy(ids).to(dev)
    prep = miner.prepare(d_off, d_ids, len(cols), 0.05)
    for s in (0.05, 0.07, 0.1):
        shared = miner.mine_csr(d_off, d_ids, len(cols), s, prepared=prep).as_dict()
        fresh = miner.mine_csr(d_off, d_ids, len(cols), s).as_dict()
        assert shared == fresh
    with pytest.raises(ValueError):
        miner.mine_csr(d_off, d_ids, len(cols), 0.01, prepared=prep)


@pytest.mark.gpu
def test_host_tables_stream_equals_device_tables():
    off, items = synth.zipf_playlists(5000, 400, mean_draws=25, alpha=0.9, seed=8)
    ht = api.HostTables(torch.device("cuda", 0))
    for _ in range(2):                      # second use reuses the pinned buffers
        res, rules = kmls_b200.mine_ids(torch.from_numpy(off).pin_memory(), torch.from_numpy(items).pin_memory(),
                                        400, 0.02, 0.5, host_tables=ht)
        assert len(ht.levels) == len(res.levels)
        assert (ht.row_items == res.row_items.cpu().numpy()).all()
        for (hi, hc), (di, dc) in zip(ht.levels, res.levels):
            assert (hi == di.cpu().numpy()).all() and (hc == dc.cpu().numpy()).all()
