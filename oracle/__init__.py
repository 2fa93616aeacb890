"""CPU oracle for the GNN-UDS Emulator hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``gnn_uds_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and only as the checker or
as the timed CPU baseline - never as the thing shipped.

PARITY UNPINNED: the reference (Zhiyu014/GNN-UDS) ships no tests, golden
vectors or trained weights, and its arithmetic lives in third-party packages
that are not installable here (tensorflow 2.10.0, keras 2.10.0, spektral 1.3.1,
swmm-api 0.2.0.18.3; reference ``requirements.txt:1-9``).  This oracle restates
their published layer semantics (Dense, Conv1D causal/dilated, GRU
reset_after=True, LSTM, spektral GCNConv/GATConv dense mode, gcn_filter) and
follows the reference's own call sites line by line; each function cites the
reference file:line it follows.
"""
