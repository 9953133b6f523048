"""Seeded random differential test: the plain-C restatement (oracle/vkt_oracle.c) against the UNMODIFIED reference compiled
from its own sources (oracle/_ref/libvkt_ref.so) on about 2600 small random cases per run -- random dims, formats,
mappings, ranges, offsets, bin counts, halos (tests/fuzz_sweeps.py).  The fixed cases of tests/test_oracle_vs_ref.py pin
the documented corner cases; this sweep looks for the ones nobody thought of.  Ranges stay inside the DESTINATION (the
reference does not check them and would scribble over the heap); CopyRange / BrickDecompose sources may leave the volume
(clamped reads).  Deterministic: numpy RandomState with fixed seeds; a few seconds."""
import pytest

import fuzz_sweeps as fz
from oracle import pyoracle as po

pytestmark = pytest.mark.needs_ref


@pytest.mark.parametrize("sweep,n", [(fz.sweep_fill_transform, 400), (fz.sweep_copy_range, 500), (fz.sweep_arithmetic, 600),
                                     (fz.sweep_resample, 400), (fz.sweep_resample_aligned, 25), (fz.sweep_flip_scan_histogram_bricks, 400),
                                     (fz.sweep_scale_rotate_filter, 300), (fz.sweep_aggregates, 500), (fz.sweep_transform2, 300)], ids=lambda v: getattr(v, "__name__", str(v)))
def test_restatement_equals_the_compiled_reference(sweep, n):
    sweep(po.oracle(), po.reference(), n, True)
