"""Golden response of the reference's own FastAPI app (/root/reference/app/main.py, unmodified, on the Keras
shim) for one synthetic request.  Build container only:  python tests/golden/make_api_golden.py"""
import base64
import contextlib
import io
import os
import sys
from urllib.parse import unquote

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import reference_loader  # noqa: E402
from deconv_api_b200 import synth  # noqa: E402


def main():
    from fastapi.testclient import TestClient
    ref_main = reference_loader.load_main()
    client = TestClient(ref_main.app)
    img = synth.smooth_image_u8(21, 200, 300)
    ok, png = cv2.imencode(".png", img)
    uri = "data:image/png;base64," + base64.b64encode(png.tobytes()).decode("ascii")
    out = {"request_png": np.frombuffer(png.tobytes(), np.uint8)}
    for layer in ("block3_conv3", "block1_conv2"):
        with contextlib.redirect_stdout(io.StringIO()):
            r = client.post("/", data={"file": uri, "layer": layer})
        assert r.status_code == 200, r.text
        body = r.json()
        assert body.startswith("data:image/webp;base64,")
        jpg = base64.b64decode(unquote(body.split(",", 1)[1]))
        dec = cv2.imdecode(np.frombuffer(jpg, np.uint8), cv2.IMREAD_COLOR)
        assert dec.shape == (448, 448, 3)
        out["jpeg|" + layer] = np.frombuffer(jpg, np.uint8)
    r = client.get("/health-check")
    out["health"] = np.frombuffer(r.content, np.uint8)
    np.savez_compressed(os.path.join(HERE, "api_reference.npz"), **out)
    print("api_reference.npz %.1f KB" % (os.path.getsize(os.path.join(HERE, "api_reference.npz")) / 1024.0))


if __name__ == "__main__":
    main()
