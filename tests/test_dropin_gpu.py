"""Drop-in boundary on the GPU: ONE object file, compiled from tests/dropin/mpc_user.c against the
reference's own headers, linked once against the unmodified reference and once against
librldl_b200.so.  Both programs must print the same closed-loop MPC run (cold solve, two plant steps
with osqp_partial_update_bounds + osqp_warm_start, osqp_update_lin_cost, osqp_update_recursive shrink and
grow, osqp_update_bounds + tighter tolerances, a refused invalid bound update): same return codes,
dimensions, status and iteration count at every solve, x / y / objective within 1e-6 relative
(north_star).  The golden fixture is the reference-linked program's output committed from the build
container; on a box that carries oracle/_ref the reference-linked program is also run live."""
import os

import pytest

from dropin_common import GOLDEN, GOLDEN_API, assert_same_run, binary, parse, run

pytestmark = pytest.mark.gpu
REL_TOL = 1e-6


@pytest.mark.parametrize("name,golden,ref_name", [("dropin_b200", GOLDEN, "dropin_ref"),
                                                  ("dropin_b200_ownhdr", GOLDEN, "dropin_ref"),
                                                  ("api_b200", GOLDEN_API, "api_ref")])
def test_user_program_linked_against_the_product_matches_the_reference(name, golden, ref_name):
    """api_b200 (tests/dropin/api_user.c): the stock API on the reference's basic_qp problem (osqp_setup,
    osqp_solve, update_lin_cost, update_lower/upper_bound, warm_start_x/_y, update_rho, update_P_A,
    update_max_iter, cleanup), the combine variant (setup 15+3, solve, update_Z_horizon +2), the reference's small
    test problems, and part D: osqp_set_default_settings as it is (Ruiz scaling 10, adaptive rho) + polish on
    basic_qp and on a badly scaled QP, with osqp_update_bounds / osqp_update_P / osqp_update_polish in between."""
    assert os.path.exists(binary(name)), f"{name} not built: run `python __graft_entry__.py`"
    cp = run(name)
    assert cp.returncode == 0, (cp.returncode, cp.stdout[-500:], cp.stderr[-2000:])
    got = parse(cp.stdout)
    assert_same_run(got, parse(open(golden).read()), REL_TOL, f"{name} vs golden")
    if os.path.exists(binary(ref_name)):
        ref = run(ref_name)
        assert ref.returncode == 0, ref.stderr[-2000:]
        assert_same_run(got, parse(ref.stdout), REL_TOL, f"{name} vs live reference")


@pytest.mark.parametrize("batch,threads", [(2048, 3), (64, 4), (4096, 8)])
def test_c_host_sharding_over_threads_and_devices_agrees_with_one_handle(batch, threads):
    """SURVEY §8e from a C host (tests/dropin/sharded_user.c): contiguous shards, one host thread and one
    RLDLBatch handle per shard, thread t on device t % G, no collective; same status / iterations / x / y as
    the whole batch through one handle, for a cold solve and a shifted warm-started closed-loop step."""
    import subprocess
    exe = binary("sharded_b200")
    assert os.path.exists(exe), "sharded_b200 not built: run `python __graft_entry__.py`"
    cp = subprocess.run([exe, str(batch), str(threads)], capture_output=True, text=True, timeout=600)
    assert cp.returncode == 0, (cp.returncode, cp.stdout[-1500:], cp.stderr[-1500:])
    assert "SHARDED OK" in cp.stdout and cp.stdout.count("AGREE solve") == 2
    assert cp.stdout.count("SHARD ") >= threads
