"""Elastic membership end to end on CPU: real rank processes (dry-run trainer)
under the supervisor, the reconciler scaling pods and syncing np, the
etcd-compatible store over HTTP.  A rescale must not lose the step."""
import tempfile

from paddle_operator_b200.workloads import elastic_demo


def test_rescale_4_2_4_keeps_the_step():
    work = tempfile.mkdtemp(prefix="pdj-elastic-test-")
    res = elastic_demo.run([4, 2, 4], steps_per_phase=20, dry=True, batch=1, tiny=True, gpus=[0, 1, 2, 3],
                           work_root=work, timeout=90)
    assert [p["world"] for p in res["phases"]] == [4, 2, 4]
    assert [p["gen"] for p in res["phases"]] == [0, 1, 2]
    assert len(res["rescales"]) == 2
    for r in res["rescales"]:
        assert r["resume_step"] == r["stop_step"], r  # the new generation continues at the very step the old one stopped at
        assert r["recovery_seconds"] < 30
    # the dry-run "model state" is a counter incremented once per step: after hand-over it must equal the step
    for p in res["phases"]:
        assert abs(p["loss"] - p["step"]) < 1e-6 or p["loss"] <= p["step"], p
    assert res["events"] == ["Scaled: scaled replicas to 2", "Scaled: scaled replicas to 4"]
    # the scale-up adopted the two parked ranks the supervisor had started after the scale-down
    assert res["rescales"][1]["adopted_standbys"] == 2, res["rescales"]
    assert res["job_phase"] in ("Running", "Starting", "Pending")


def test_killed_rank_is_restarted_and_state_is_handed_over():
    """SIGKILL one rank mid-run (dry-run trainers do not block on a collective, so the survivors notice the
    expired lease, stop together, and re-form with the restarted rank without going back to the checkpoint)."""
    work = tempfile.mkdtemp(prefix="pdj-elastic-fail-")
    res = elastic_demo.run_failure(3, fail_rank=1, fail_step=15, batch=1, tiny=True, gpus=[0, 1, 2], work_root=work,
                                   timeout=90, dry=True)
    assert res["resumed_generation"] >= 1 and res["rank_restarts"] >= 1
    assert res["steps_redone"] <= 1 and res["recovery_seconds"] < 30
