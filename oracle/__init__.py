"""CPU oracle for the bundle-adjustment hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import this
package; ``mpsfm_b200`` never does (a product path routed through here would void every parity claim).

What it restates (plain numpy/scipy, float64):

* the residual blocks the reference hands to Ceres in
  /root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:104 (COLMAP reprojection error, all camera
  models) and :163-176 (fork-only log-depth prior),
* Ceres 2.1's loss functions, corrector, ``EigenQuaternionManifold``/``SubsetManifold``,
* Ceres 2.1's ``TrustRegionMinimizer`` + ``LevenbergMarquardtStrategy`` with an exact Schur
  complement solve (``SPARSE_SCHUR``, bundle_adjustment.py:285-293).

PARITY UNPINNED: the arithmetic lives in third-party code that is not under /root/reference and is
not installable here -- COLMAP fork ``github.com/Zador-Pataki/colmap`` (unpinned HEAD,
docker/install_colmap.sh:28), ceres-solver 2.1.0 (docker/install_ceres_solver.sh:12), pyceres
(docker/install_pyceres.sh:2-4) -- and the reference ships no tests, golden vectors or logged
solver output for this path (SURVEY.md sections 4 and 8c).  This oracle restates the published
algorithms of those libraries and the reference's own call sites; it is pinned only against
independent derivations (forward-mode duals vs. finite differences / sympy, dense normal equations
vs. Schur elimination), not against outputs of the real reference.
"""
