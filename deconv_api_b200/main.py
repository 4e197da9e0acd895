"""Drop-in for reference ``app/main.py``: the same FastAPI application — routes, form fields, CORS policy,
response format — with the visualisation on the B200 path.

    GET  /health-check          -> {"healthy": "true"}                                  (main.py:41-43)
    POST /  (form: file, layer) -> JSON string 'data:image/webp;base64,' + quote(b64(JPEG))  (main.py:45-78)

Run: ``uvicorn deconv_api_b200.main:app --port 80`` (reference Dockerfile:15 runs ``uvicorn main:app``).
Image decode / resize / JPEG encode stay on the CPU (OpenCV), exactly as in the reference; everything between
``preprocess_input`` and the uint8 mosaic runs through libdcv.so.
"""
import base64
import os
from base64 import b64encode
from urllib.parse import quote

import cv2
import numpy as np
from fastapi import FastAPI, Form
from fastapi.middleware.cors import CORSMiddleware

from . import vgg16
from .deepdream import deprocess_image, visualize_all_layers
from .vgg16 import preprocess_input

# main.py:16-17.  ImageNet weights cannot be downloaded here; DCV_WEIGHTS may point to an .npz export.
model = vgg16.VGG16(weights=os.environ.get("DCV_WEIGHTS", "imagenet"), include_top=True)

# The reference projects the top-8 maps of EVERY layer below the requested one (136 projections for
# block5_conv3) and then uses four of one layer (main.py:67-68).  Default = that behaviour, step for step.
# DCV_API_FAST=1 selects the fused request tail instead (one C-ABI call: preprocess, forward, top-4 of the
# requested layer only, their projections, mosaic and de-process on the GPU; 150 KB up, 602 KB down) — the
# response is the same image.
FAST_TAIL = os.environ.get("DCV_API_FAST", "0") == "1"
PRECISION = os.environ.get("DCV_API_PRECISION", "fp32")

app = FastAPI()

origins = [
    "*",
]

app.add_middleware(
    CORSMiddleware,
    allow_origins=origins,
    allow_credentials=False,
    allow_methods=["*"],
    allow_headers=["*"],
)


def readb64(uri):
    # main.py:35-39: text after the first comma is base64 of any OpenCV-decodable image; BGR uint8 out
    encoded_data = uri.split(',')[1]
    nparr = np.frombuffer(base64.b64decode(encoded_data), np.uint8)
    img = cv2.imdecode(nparr, cv2.IMREAD_COLOR)
    return img


@app.get('/health-check')
def healthcheck():
    return {'healthy': 'true'}


@app.post('/')
async def return_deconv(file: str = Form(...), layer: str = Form(...)):
    img = readb64(file)
    img = cv2.resize(img, (224, 224))                       # main.py:53 (bilinear)
    layer_name = layer

    if FAST_TAIL:
        total_concat = mosaic_on_gpu(img, layer_name)
        return encode_response(total_concat)

    # main.py:60-61: img_to_array -> expand_dims -> preprocess_input ('caffe'); the decoded image is BGR and
    # gets flipped again, so the channel means are subtracted in the reference's (mismatched) order on purpose
    in_array = np.expand_dims(np.asarray(img, dtype=np.float32), axis=0)
    img_array = preprocess_input(in_array)

    deconv = visualize_all_layers(model, img_array, layer_name=layer_name, visualize_mode='all', precision=PRECISION)

    # main.py:67-69: 2x2 mosaic of the four strongest maps (KeyError / IndexError propagate -> HTTP 500)
    top_img = np.concatenate((deconv[layer_name][0], deconv[layer_name][1]), axis=1)
    bottom_img = np.concatenate((deconv[layer_name][2], deconv[layer_name][3]), axis=1)
    concat_img = np.concatenate((top_img, bottom_img), axis=0)

    total_concat = deprocess_image(concat_img)
    return encode_response(total_concat)


def encode_response(total_concat):
    # main.py:73-78: JPEG (cv2 default quality), base64, percent-quoted, behind a 'webp' data-URL prefix
    retval, buffer = cv2.imencode('.jpg', total_concat)
    data = b64encode(buffer)
    data = data.decode('ascii')
    data_url = 'data:image/webp;base64,{}'.format(quote(data))
    return data_url


def mosaic_on_gpu(img_bgr_u8, layer_name):
    """Fused request tail (dcv_request_mosaic).  Keeps the reference's error behaviour: an unknown layer name makes
    visualize_all_layers stack every layer and main.py:67 then raises KeyError; fewer than four positive maps is an
    IndexError (main.py:67-68)."""
    from .deepdream import adopt
    m = adopt(model)
    if layer_name not in [l.name for l in m.layers][1:]:
        raise KeyError(layer_name)
    eng = m.bound_engine(None, upto=layer_name)
    mosaic, _ = eng.request_mosaic(img_bgr_u8, eng.index[layer_name], mode='all', precision=PRECISION)
    return mosaic
