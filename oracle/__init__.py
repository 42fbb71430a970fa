"""CPU oracle for the CC-MPC prediction -> chance-constraint hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it, and there only as the
checker (or as the timed CPU baseline), never as the thing shipped.  The
product path (``ccmpc_b200``) never imports this package and fails loudly when
its CUDA library is missing.

What is in here
---------------
``restate.py``          NumPy/SciPy restatement of the reference algorithm, function by function, each citing the
                        reference ``file:line`` it follows (all paths relative to ``/root/reference``).
``restate_planner.py``  The planner-level drivers around it: the Minkowski / affine builders with their
                        shrinking-horizon branch, save_moments hand-off, L4 polytopes, bookkeeping tuples.
``ref_loader.py``       Imports the reference's own ``makeconstraint.py`` and ``utility/npu.py`` *by file path*, and
                        ``extract()``: compiles single functions / methods out of reference files whose MODULES do
                        not import here (carla, cvxpy, docplex, matplotlib, control, openpyxl at module level) but
                        whose bodies use only numpy / scipy / pickle / os / logging.  This container only - never on
                        the GPU box.
``ref_planner.py``      Runs the reference's own ``MidlevelAgent`` hot-path methods (extracted, unmodified) on an
                        object whose attributes are set by hand; cvxpy and ``np.random`` are shimmed (rows are
                        recorded; noise comes from a seeded legacy ``RandomState``).
``make_golden.py``, ``make_golden_extracted.py``, ``make_golden_planner.py``
                        Run the reference's own code on seeded inputs and write ``tests/golden/*.npz`` (committed).
``synth.py``            Re-export of ``ccmpc_b200.synth`` (seeded input generators; they moved out of this package
                        because the product arm of bench.py must import nothing from here).

Parity pinning
--------------
The reference's tests hold NO golden vector on this path, so the oracle is pinned against outputs of
the reference's own code run in this container (``tests/golden/ref_*.npz``; ``tests/test_oracle_golden.py`` and
``tests/test_oracle_planner_golden.py`` hold the restatement to them, the ``-m gpu`` tests hold the kernels to them):

  makeconstraint.py (imported)        compute_mvoe, predict_moments, tangent_lines_of_slope_m,
                                      distance_point_to_line, choose_closest_tangent, compute_scale,
                                      compute_lower_bound                                  ref_mvoe / _predict_moments /
                                                                                           _tangent / _scale_lower / _pipeline
  utility/npu.py (imported)           vertices_of_bboxes                                   ref_vertices
  midlevel/util.py:171-200            compute_L4_outerapproximation (extracted)            ref_l4
  v8ideal/__init__.py:2575-2618       MidlevelAgent.save_moments (extracted)               ref_moments
  v8ideal/__init__.py:2620-2711       MidlevelAgent.predict_ideal (extracted; its 1e6-sample noise is replayed
                                      from the seed, the rows it drew are stored)          ref_ideal_rollout
  midlevel/ovehicle.py:23-117         OVehicle.from_trajectron (extracted)                 ref_grouping
  dataForCost_ref.py:198-232          getData_singleEpisode.cost_final (extracted)         ref_cost
  v8ideal/__init__.py:781-964         compute_obstacle_constraints_GMM_Minkowski_idealprediction (extracted), with
  v8ideal/__init__.py:1378-1539       compute_obstacle_constraints_GMM_affine, __compute_vertices,
  v8ideal/__init__.py:627-640,        __compute_overapproximations: three consecutive shrinking-horizon steps at c1
                      694-736         (moments handed over through the reference's own pickle), K growing between
                                      steps, the c2 step                                   ref_planner_c1 / _kgrow / _c2

PARITY UNPINNED (nothing in the reference to run):
  * the sample-level collision / violation counts (A9): NOT IN THE REFERENCE, restated from its constants
    (R = 3.4, v8ideal/__init__.py:795);
  * GMM2D sampling and unicycle / single-integrator integration: the code is in the Trajectron++ submodule
    (fireofearth/Trajectron-plus-plus @ 0e44e82), which is absent from /root/reference; restated from the published
    model (``restate.gmm2d_sample``, ``restate.integrate_samples``).
"""
