"""Keras .h5 subset reader/writer (gnn_uds_b200/h5lite.py): round trips and name mapping.
No h5py/libhdf5 in this image, so a real Keras file cannot be checked here."""
import os
import struct

import numpy as np
import pytest

import cases
from gnn_uds_b200 import h5lite, weights as Wt
from oracle.emulator_oracle import OracleEmulator


def _specs(case):
    args, B, seed = cases.build_args(case)
    o = OracleEmulator(args.conv, args.resnet, args.recurrent, args, dtype=np.float32)
    shim = cases.SpecShim(o)
    specs = Wt.layer_specs(shim)
    return specs, Wt.init_weights(specs, seed, node_edge_bias_std=0.05, bias_std=0.05)


@pytest.mark.parametrize("case", ["c2_shq_gat_conv_10", "c1_ast_gcn_gru", "ast_gcn_lstm_act_nofusion"])
def test_round_trip(tmp_path, case):
    specs, ws = _specs(case)
    p = str(tmp_path / "model.h5")
    h5lite.save_keras_weights(p, specs, ws, extra_layers=["input_1", "tf.reshape_3", "tf.__operators__.getitem"])
    got = h5lite.load_keras_weights(p, specs)
    assert len(got) == len(ws)
    for a, b in zip(got, ws):
        for x, y in zip(a, b):
            assert x.dtype == np.float32 and np.array_equal(x, y)


def test_keras_names_and_paths(tmp_path):
    specs, ws = _specs("c1_ast_gcn_gru")
    names = Wt.keras_names(specs)
    assert names[:3] == ["dense", "dense_1", "dense_2"] and "dense_resx" in names and "gru_3" in names
    p = str(tmp_path / "model.h5")
    h5lite.save_keras_weights(p, specs, ws)
    layers = dict(h5lite.read_keras_layers(p))
    assert [w for w, _ in layers["gru_1"]] == ["gru_1/gru_cell_1/kernel:0", "gru_1/gru_cell_1/recurrent_kernel:0",
                                               "gru_1/gru_cell_1/bias:0"]
    assert [w for w, _ in layers["gcn_conv"]] == ["gcn_conv/kernel:0", "gcn_conv/bias:0"]
    f = h5lite.H5File(p)
    assert f.root["gru_1/gru_1/gru_cell_1/kernel:0"].read().shape == (64, 192)
    assert "dense_resx" in f.root


def test_offset_auto_names_load_by_order(tmp_path):
    """A file saved from a session whose name counters were not fresh (mbrl.py builds agents first)."""
    specs, ws = _specs("c3_ast_act_5")
    p = str(tmp_path / "model.h5")
    h5lite.save_keras_weights(p, specs, ws, name_start={"dense": 11, "mixed_gat": 4, "conv1d": 2, "node_edge": 6})
    layers = [n for n, _ in h5lite.read_keras_layers(p)]
    assert "dense_11" in layers and "dense" not in layers
    got = h5lite.load_keras_weights(p, specs)
    for a, b in zip(got, ws):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)


def test_shape_mismatch_and_missing_layers_fail_loudly(tmp_path):
    specs, ws = _specs("c1_ast_gcn_gru")
    p = str(tmp_path / "model.h5")
    h5lite.save_keras_weights(p, specs, ws)
    other, _ = _specs("ast_gat_gru_noflood")
    with pytest.raises(h5lite.H5Error):
        h5lite.load_keras_weights(p, other)
    with pytest.raises(h5lite.H5Error):
        h5lite.load_keras_weights(p, specs[:-1])      # file has one weighted layer too many


def test_not_hdf5(tmp_path):
    p = tmp_path / "junk.h5"
    p.write_bytes(b"\0" * 4096)
    with pytest.raises(h5lite.H5Error):
        h5lite.H5File(str(p))


def test_continuation_block_and_v3_attribute(tmp_path):
    """Reader-side features the writer does not emit: object-header continuation (0x0010) and
    attribute message version 3 (what h5py writes for UTF-8 named attributes)."""
    w = h5lite._Writer()
    w.leaf_k = 4
    dt, ds = w._dt_str(5), w._dataspace((2,))
    name = b"weight_names\0"
    body = struct.pack("<BBHHHB", 3, 0, len(name), len(dt), len(ds), 1) + name + dt + ds + b"a:0\0\0b:0\0\0"
    cont = w._msg(0x000C, body)
    caddr = w.alloc(cont)
    kids = {"a:0": w.dataset(np.arange(3, dtype=np.float32)), "b:0": w.dataset(np.ones((2, 2), np.float32))}
    names = sorted(kids)
    heap = bytearray(8); offs = {}
    for nm in names:
        offs[nm] = len(heap); b = nm.encode() + b"\0"; heap += b + b"\0" * ((-len(b)) % 8)
    dseg = w.alloc(bytes(heap))
    haddr = w.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), 1, dseg))
    ent = b"".join(struct.pack("<QQII16x", offs[nm], kids[nm], 0, 0) for nm in names) + b"\0" * (40 * 6)
    snod = w.alloc(b"SNOD" + struct.pack("<BBH", 1, 0, 2) + ent)
    tree = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, h5lite.UNDEF, h5lite.UNDEF) + struct.pack("<QQQ", 0, snod, offs[names[-1]])
    baddr = w.alloc(tree + b"\0" * 512)
    hdr = w.header([w._msg(0x0011, struct.pack("<QQ", baddr, haddr)), w._msg(0x0010, struct.pack("<QQ", caddr, len(cont)))])
    # the header's message count must include the continued message
    struct.pack_into("<H", w.buf, hdr + 2, 3)
    p = tmp_path / "c.h5"
    p.write_bytes(w.finish(hdr, baddr, haddr))
    f = h5lite.H5File(str(p))
    assert [s.decode() for s in f.root.attrs()["weight_names"]] == ["a:0", "b:0"]
    assert np.array_equal(f.root["a:0"].read(), np.arange(3, dtype=np.float32))
