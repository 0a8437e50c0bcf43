"""oracle/pipeline_ref.py against the committed golden vectors, i.e. against the reference's own
model/*.py executed over the oracle (generator: oracle/make_golden.py)."""
import numpy as np
import pytest

from oracle import pipeline_ref, rans_ref
from tests.helpers import align_rows, golden_config, load_golden


@pytest.mark.parametrize("name", ["ref_inverse_scaling", "ref_inverse_nn", "ref_fixed_R1"])
def test_pipeline_ref_reproduces_reference_run(name):
    z, sd = load_golden(name)
    cfg = golden_config(z)
    codec = pipeline_ref.RefCodec(sd, cfg)
    q = z["q"].tolist()
    enc = codec.encode_block(z["points"], z["colors"], q)
    assert [v[0] for v in enc["k"]] == z["k"].tolist()
    assert (enc["y_points"] == z["y_coords"][np.argsort(rows_key(z["y_coords"]))]).all()
    ys = rans_ref.encode_with_indexes(enc["y_symbols"].reshape(-1), enc["y_indexes"].reshape(-1),
                                      codec.gc._quantized_cdf.numpy(), codec.gc._cdf_length.numpy(), codec.gc._offset.numpy())
    assert ys == z["y_string"].tobytes()
    C = enc["z_symbols"].shape[1]
    zs = rans_ref.encode_with_indexes(enc["z_symbols"].reshape(-1), np.repeat(np.arange(C), enc["z_symbols"].shape[2]),
                                      codec.eb._quantized_cdf.numpy(), codec.eb._cdf_length.numpy(), codec.eb._offset.numpy())
    assert zs == z["z_string"].tobytes()
    rec = codec.decode_block(enc["y_points"], enc["z_symbols"], enc["y_symbols"], enc["k"], q)
    ref = z["reconstruction"]
    assert rec.shape == ref.shape
    ia, ib = align_rows(rec[:, :3], ref[:, :3])
    assert len(ia) == len(ref), "decoded geometry differs from the reference run"
    assert np.abs(rec[ia, 3:] - ref[ib, 3:]).max() < 1e-6


def rows_key(C):
    from oracle import me_ref
    return me_ref.pack_keys(C)
