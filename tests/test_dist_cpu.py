"""N > 1 host path on the CPU: two gloo ranks, members placed i mod world, the evaluation-score all-gather as the only
collective, identical successive-halving decisions on every rank."""
import json
import os
import socket
import subprocess
import sys


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run_ranks(tmp_path, mode="plain", world=2):
    port = _free_port()
    worker = os.path.join(os.path.dirname(__file__), "dist_worker.py")
    procs = [subprocess.Popen([sys.executable, worker, str(r), str(world), str(port), str(tmp_path), mode])
             for r in range(world)]
    for p in procs:
        assert p.wait(timeout=240) == 0
    return tuple(json.load(open(tmp_path / f"rank{r}.json")) for r in range(world))


def _run_two_ranks(tmp_path, mode="plain"):
    return _run_ranks(tmp_path, mode, 2)


def test_two_rank_save_then_resume_matches_uninterrupted_run(tmp_path):
    """Every rank writes its own shard (its members' state + the replicated bookkeeping: slot order, remote step
    counts) and a job resumed from the shards ends exactly where the uninterrupted one does."""
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    plain = _run_two_ranks(tmp_path / "a")
    resumed = _run_two_ranks(tmp_path / "b", "resume")
    for p, r in zip(plain, resumed):
        assert p["candidates"] == r["candidates"]
        assert p["steps"] == r["steps"]
        assert p["scores"] == r["scores"]


def test_two_rank_trainer_agrees_on_pruning(tmp_path):
    r0, r1 = _run_two_ranks(tmp_path)
    assert r0["local_members"] == 4 and r1["local_members"] == 4          # member i -> rank i mod 2
    assert r0["candidates"] == r1["candidates"] and len(r0["candidates"]) == 2
    assert r0["scores"] == r1["scores"] and len(r0["scores"]) == 8        # every rank saw every score
    assert not set(r0["steps"]) & set(r1["steps"])                          # no member trained twice
    trained = {**r0["steps"], **r1["steps"]}
    assert sorted(trained.values()) == [10, 10, 10, 10, 20, 20, 40, 40]
    # same decisions as the single-process run
    best = sorted((float(a) for a in trained), key=lambda a: abs(__import__("math").log10(a) - 2))[:2]
    assert sorted(best) == r0["candidates"]


def test_two_rank_survivor_migration_changes_placement_not_results(tmp_path):
    """The four best candidates all start on rank 0. Without migration rank 1 idles from the second rung on; with it the
    survivors are spread again (two per rank, then one per rank), every member keeps training from the state it had --
    the device step counts travel with it -- and scores, survivors and step totals equal the static run's."""
    (tmp_path / "s").mkdir(); (tmp_path / "m").mkdir(); (tmp_path / "r").mkdir()
    static = _run_two_ranks(tmp_path / "s", "static")
    moved = _run_two_ranks(tmp_path / "m", "migrate")
    assert static[0]["migrations"] == 0 and moved[0]["migrations"] == moved[1]["migrations"] >= 2
    assert sorted(static[0]["steps"]) == sorted(["100.0", "90.0", "120.0", "70.0"])     # the premise: all on rank 0
    assert moved[0]["owner"] == moved[1]["owner"]                                        # replicated bookkeeping
    assert moved[1]["imports"] and not static[1]["imports"]
    for run in (static, moved):
        assert run[0]["candidates"] == run[1]["candidates"] and run[0]["scores"] == run[1]["scores"]
        assert not set(run[0]["steps"]) & set(run[1]["steps"])                           # one home per member
    assert moved[0]["candidates"] == static[0]["candidates"]
    assert moved[0]["scores"] == static[0]["scores"]
    assert {**moved[0]["steps"], **moved[1]["steps"]} == {**static[0]["steps"], **static[1]["steps"]}
    # the migrants' device state came along: what the population holds equals what the host believes
    for r in moved:
        assert r["device_steps"] == r["steps"]
    # survivors end up one per rank
    final = [a for a in moved[0]["candidates"]]
    assert len(final) == 2 and {moved[0]["owner"][str(a)] for a in final} == {0, 1}
    # save-then-resume across a migration ends where the uninterrupted run does
    resumed = _run_two_ranks(tmp_path / "r", "migrate_resume")
    for m, r in zip(moved, resumed):
        assert m["candidates"] == r["candidates"] and m["steps"] == r["steps"] and m["scores"] == r["scores"]
        assert m["owner"] == r["owner"] and r["device_steps"] == r["steps"]


def test_three_rank_migration_under_churning_scores_keeps_the_books_straight(tmp_path):
    """24 members on 3 ranks, successive halving by quarters, scores that reshuffle every evaluation (survivors change, pruned members re-enter), two spare
    slots per rank (a rank that has used them takes no more migrants). Whatever the plans were: every rank holds the same owner map, every member
    lives in exactly one slot of exactly one rank, the free-slot counts every rank keeps for the others are the truth, the
    members' device state (steps trained) followed them, and scores / survivors / step totals equal the static run's."""
    (tmp_path / "s").mkdir(); (tmp_path / "m").mkdir()
    static = _run_ranks(tmp_path / "s", "churn_static", 3)
    moved = _run_ranks(tmp_path / "m", "churn_migrate", 3)
    assert moved[0]["migrations"] >= 3 and all(r["migrations"] == moved[0]["migrations"] for r in moved)
    assert all(r["owner"] == moved[0]["owner"] and r["free_count"] == moved[0]["free_count"] for r in moved)
    seen = {}
    for rank, r in enumerate(moved):
        assert set(r["steps"]) == {a for a, o in r["owner"].items() if o == rank} == set(r["slots"])
        slots = sorted(r["slots"].values())
        assert len(set(slots)) == len(slots) and all(0 <= s < r["capacity"][rank] for s in slots)
        assert sorted(slots + r["free_slots"]) == list(range(r["capacity"][rank]))
        assert r["free_count"][rank] == len(r["free_slots"])
        assert r["device_steps"] == r["steps"]
        for a in r["steps"]:
            assert a not in seen
            seen[a] = rank
    assert len(seen) == 24
    assert moved[0]["candidates"] == static[0]["candidates"] and moved[0]["scores"] == static[0]["scores"]
    merged = lambda run: {a: n for r in run for a, n in r["steps"].items()}
    assert merged(moved) == merged(static)
