"""End-to-end request through the FastAPI mirror on the GPU against the response of the reference's own
app/main.py (run unmodified on the Keras shim, tests/golden/make_api_golden.py)."""
import base64
from urllib.parse import unquote

import cv2
import numpy as np
import pytest
from fastapi.testclient import TestClient

pytestmark = pytest.mark.gpu


def _post(client, png, layer):
    uri = "data:image/png;base64," + base64.b64encode(png).decode("ascii")
    return client.post("/", data={"file": uri, "layer": layer})


@pytest.mark.parametrize("fast_tail", [False, True])
def test_post_matches_reference_response(golden, monkeypatch, fast_tail):
    """fast_tail False: the reference's own sequence (full sweep, host mosaic, deprocess_image); True: the fused
    request tail (DCV_API_FAST=1, one dcv_request_mosaic call).  Both must give the reference's response."""
    from deconv_api_b200 import main as api
    monkeypatch.setattr(api, "FAST_TAIL", fast_tail)
    g = golden("api_reference.npz")
    client = TestClient(api.app)
    for layer in ("block3_conv3", "block1_conv2"):
        r = _post(client, g["request_png"].tobytes(), layer)
        assert r.status_code == 200 and r.headers["content-type"].startswith("application/json")
        body = r.json()
        assert isinstance(body, str) and body.startswith("data:image/webp;base64,")
        payload = body.split(",", 1)[1]
        assert "+" not in payload and "=" not in payload          # urllib.parse.quote: '+' -> %2B, '=' -> %3D, '/' kept
        jpg = base64.b64decode(unquote(payload))
        assert jpg[:3] == b"\xff\xd8\xff"
        got = cv2.imdecode(np.frombuffer(jpg, np.uint8), cv2.IMREAD_COLOR).astype(np.int16)
        ref = cv2.imdecode(g["jpeg|" + layer], cv2.IMREAD_COLOR).astype(np.int16)
        assert got.shape == ref.shape == (448, 448, 3)
        d = np.abs(got - ref)
        # the uint8 mosaics differ by at most one level on a few pixels (fp32 summation order); after JPEG
        # quantisation that can move an 8x8 block by a few levels
        assert d.mean() < 0.25 and np.percentile(d, 99.9) <= 6, (layer, d.mean(), d.max())


@pytest.mark.parametrize("fast_tail", [False, True])
def test_unknown_layer_is_500_like_reference(golden, monkeypatch, fast_tail):
    from deconv_api_b200 import main as api
    monkeypatch.setattr(api, "FAST_TAIL", fast_tail)
    g = golden("api_reference.npz")
    client = TestClient(api.app, raise_server_exceptions=False)
    assert _post(client, g["request_png"].tobytes(), "nope").status_code == 500
