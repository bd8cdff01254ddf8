"""CPU oracle for the rigid resample + SSD / joint-histogram-MI hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``medicalimageanalysis_b200/`` may import this package;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may (as the checker / the timed CPU arm, never as the product).

PARITY UNPINNED.  The reference (caleb-oconnor/MedicalImageAnalysis v2.1.92) has no tests, fixtures or
golden vectors, cannot be imported here (needs gdcm / pydicom / vtk / SimpleITK), and the arithmetic of
this path does not live in its repository: it lives in the third-party wheels it pins,
  * VTK 9.3.1 (``requirements.txt:82``) -- ``vtkImageReslice`` / ``vtkImageInterpolator``
  * SimpleITK 2.4.1 / ITK 5.4 (``requirements.txt:72``) -- ``ResampleImageFilter`` +
    ``LinearInterpolateImageFunction``
neither of which is vendored or installable in this image.  This package therefore restates the
PUBLISHED algorithm of those classes, anchored on the reference's own call sites
(``structure/rigid.py:646-695``, ``structure/image.py:160-215``, ``structure/deformable.py:575-579,778-795``,
``utils/deformable/simpleitk.py:76-94``) and, independently, cross-checked on interior voxels against
``scipy.ndimage.map_coordinates(order=1)`` -- a function the reference itself calls
(``structure/deformable.py:873-878``).  The FFS permutation IS pinned: it is NumPy code in the reference
(``read/dicom.py:655-740``) and is restated with the same NumPy calls.
"""
