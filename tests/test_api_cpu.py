"""Schema of the FastAPI mirror (deconv_api_b200/main.py) against reference app/main.py — the parts that need
no GPU: routes, operation ids, form fields, 422 on a missing field, CORS, health check body."""
import numpy as np
from fastapi.testclient import TestClient

from deconv_api_b200 import main as api


def test_health_check_matches_reference(golden):
    client = TestClient(api.app)
    r = client.get("/health-check")
    assert r.status_code == 200 and r.json() == {"healthy": "true"}
    assert r.content == golden("api_reference.npz")["health"].tobytes()


def test_openapi_schema_matches_reference():
    spec = api.app.openapi()
    assert set(spec["paths"]) == {"/health-check", "/"}
    assert spec["paths"]["/health-check"]["get"]["operationId"] == "healthcheck_health_check_get"
    post = spec["paths"]["/"]["post"]
    assert post["operationId"] == "return_deconv__post"
    body = post["requestBody"]["content"]["application/x-www-form-urlencoded"]["schema"]
    ref = body["$ref"].split("/")[-1]
    props = spec["components"]["schemas"][ref]
    assert set(props["properties"]) == {"file", "layer"} and set(props["required"]) == {"file", "layer"}


def test_missing_field_is_422_and_cors_is_open():
    client = TestClient(api.app)
    assert client.post("/", data={"layer": "block1_conv1"}).status_code == 422
    assert client.post("/", data={"file": "x,y"}).status_code == 422
    r = client.get("/health-check", headers={"Origin": "http://example.org"})
    assert r.headers.get("access-control-allow-origin") == "*"
    pre = client.options("/", headers={"Origin": "http://example.org", "Access-Control-Request-Method": "POST",
                                       "Access-Control-Request-Headers": "x-custom"})
    assert pre.status_code == 200 and "POST" in pre.headers.get("access-control-allow-methods", "")


def test_readb64_decodes_bgr(golden):
    import base64
    png = golden("api_reference.npz")["request_png"].tobytes()
    img = api.readb64("data:image/png;base64," + base64.b64encode(png).decode())
    assert img.shape == (200, 300, 3) and img.dtype == np.uint8
