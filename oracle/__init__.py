"""CPU oracle for the WSInsight patch-classification hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``wsinsight_b200/`` may import this
package.  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline``
/ ``--impl reference`` legs of ``bench.py`` may import or execute it, and there
only as the checker / timed CPU baseline, never as the product path.

The reference (``/root/reference``, huangch/wsinsight) is pure Python and its
arithmetic for this path lives in third-party libraries.  The
reference package itself cannot be imported in this image (missing skimage,
shapely, h5py, tiffslide, wsinfer_zoo ...), so each function here
restates one reference call site with the *same* third-party call where that
library is importable (cv2, PIL, torchvision, torch, scipy.ndimage -- which is
what skimage calls underneath) and with a numpy restatement otherwise.

Pinning status (see DESIGN.md "Oracle"):
  * RGB->S, median, threshold: pinned against live cv2 (exhaustive 2^24 RGB).
  * closing / remove_small_objects / remove_small_holes: pinned against
    scipy.ndimage (skimage 0.25.2 itself is not importable: "unpinned vs skimage").
  * contours: live cv2.findContours; polygon algebra + ``contains``: even-odd
    restatement of Shapely/GEOS (not importable: "unpinned vs GEOS"), pinned
    against the reference's golden CSV coordinates (tests/golden/purple_coords.json).
  * resize/ToTensor/Normalize: live PIL + torchvision (these ARE the reference's
    calls) plus a fixed-point numpy restatement pinned bit-exact against PIL.
  * forward: live torchvision resnet34 fp32 / restated InceptionV4 fp32;
    probabilities of the real zoo weights are not reachable offline
    ("parity unpinned" vs the golden CSV probabilities).
"""
