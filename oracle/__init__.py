"""CPU oracle for the corrupt -> denoise preprocessing path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``image_denoising_b200/`` may import
this package; it is used by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` as the checker and
as the timed CPU baseline, never as the shipped path.

What it is: a transcription of the expressions inside the reference's closures
(``lib/roi_data_layer/minibatch.py:87-1516`` and their twins in
``lib/model/test.py:193-1607``), calling the very same third-party CPU
libraries the reference calls (OpenCV, NumPy, SciPy, scikit-learn, PIL), plus a
NumPy restatement of the two libraries that are absent from this image:
scikit-image 0.14.2 (``random_noise``, ``denoise_wavelet``, ``img_as_float``,
``rgb2ycbcr``/``ycbcr2rgb``) and PyWavelets 1.0.3 (``wavedecn``/``waverecn``/
``threshold``) - pinned in ``requirements.txt:151,160``.

Parity status: the reference ships no tests, golden vectors or fixtures for
this path, and its modules cannot be imported here
(TensorFlow 1.14 / skimage / wandb / easydict are missing).  The oracle is
therefore pinned as follows:
  * cv2 / numpy / scipy / sklearn / PIL pieces ARE the reference's own
    libraries called with the reference's own arguments - pinned by identity;
  * ``Automold.add_sun_flare`` is imported from /root/reference at fixture
    generation time (tests/golden/make_golden.py) and the outputs committed;
  * the skimage / pywt restatement is "parity unpinned" against the real
    libraries (they cannot be installed: no network); it is held by
    tests/test_oracle_wavelet.py: property tests (perfect reconstruction,
    Parseval, BT.601 anchors, hand-computed Haar cases), a second, independently
    derived transform (dense analysis / synthesis matrices from the filter
    definitions and the symmetric-extension rule) the oracle must equal on odd
    and even sizes, and the known answers printed in the PyWavelets docs;
  * ``oracle/jpeg.py`` (the ``cv2.imread`` stage in front of the path) restates libjpeg's
    baseline decoder and IS pinned: byte-identical to ``cv2.imdecode`` in tests/test_oracle_jpeg.py;
  * ``oracle/nms.py`` is pinned by golden vectors of the reference's lib/nms/py_cpu_nms.py;
  * tests/golden/manifest.json pins 525 oracle results (the whole noise x denoise x level ladder) by SHA-256 (tests/test_golden_cpu.py).
"""
from . import noise, denoise, wavelet, closures, jpeg, nms  # noqa: F401
