"""Row f3: export / import in the reference's `{cards, centroids, meta}` schema (app.mjs:263-283)."""
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("board_json", os.path.join(ROOT, "k-means-demo_b200", "board_json.py"))
bj = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bj)


def test_round_trip_is_bit_exact():
    rng = np.random.default_rng(3)
    for d, dt in ((2, np.float32), (5, np.float32), (3, np.float64)):
        x = (rng.standard_normal((40, d)) * 7).astype(dt)
        c = rng.standard_normal((4, d))
        lab = rng.integers(-1, 4, 40).astype(np.int32)
        board = bj.loads(bj.dumps(bj.export_board(x, lab, c, iteration=9, inertia=1.25)))
        back = bj.import_board(board)
        assert back["points"].dtype == dt and np.array_equal(back["points"], x)
        assert np.array_equal(back["labels"], lab) and np.array_equal(back["centroids"], c)
        assert back["iteration"] == 9 and back["inertia"] == 1.25


def test_schema_is_the_demos():
    x = np.array([[0.5, 1.0], [2.0, 3.0]], dtype=np.float32)
    b = bj.export_board(x, [1, -1], [[0.0, 0.0], [1.0, 1.0]], iteration=3)
    assert set(b) == {"cards", "centroids", "meta"}
    assert set(b["cards"][0]) == {"id", "title", "traits", "assignedTo", "createdBy"}      # app.mjs:143-145
    assert {"id", "name", "color", "locked"} <= set(b["centroids"][0])                     # app.mjs:128
    assert b["cards"][0]["assignedTo"] == "c:1" and b["cards"][1]["assignedTo"] is None
    assert b["centroids"][1]["name"] == "Centroid 2" and b["centroids"][0]["locked"] is False
    assert b["meta"]["iteration"] == 3 and b["meta"]["pos:p:0"] == {"x": 0.02, "y": 0.1}   # app.mjs:158, 288, 366-367
    assert all(isinstance(t, str) for t in b["cards"][0]["traits"])
    assert [c["count"] for c in b["centroids"]] == [0, 1]


def test_import_of_a_board_the_demo_exported():
    # the demo's own cards: word traits, centroids without coordinates (app.mjs:204-216, 128)
    board = {"cards": [{"id": "seed:t1", "title": "Nguyen", "traits": ["Sweet", "Creamy"], "assignedTo": "c:a", "createdBy": "seed"},
                       {"id": "seed:t2", "title": "Patel", "traits": ["Fresh", "Sorbet"], "assignedTo": None, "createdBy": "seed"},
                       {"id": "seed:t3", "title": "Garcia", "traits": ["Chocolatey", "Crunchy"], "assignedTo": "c:gone", "createdBy": "seed"}],
             "centroids": [{"id": "c:a", "name": "Centroid 1", "color": "#e11", "locked": False}],
             "meta": {"iteration": 2, "mode": "explore"}}
    back = bj.import_board(board)
    assert back["points"] is None and back["centroids"] is None
    assert back["labels"].tolist() == [0, -1, -1] and back["iteration"] == 2
    assert back["card_ids"] == ["seed:t1", "seed:t2", "seed:t3"]


def test_bad_input():
    with pytest.raises(ValueError):
        bj.import_board({"cards": 3})
    with pytest.raises(ValueError):
        bj.export_board(np.zeros((2, 2)), [0, 5], np.zeros((2, 2)))


def test_exported_positions_lie_in_the_demo_clamp_box():
    """`pos:<id>` is a fraction of the drop zone which the demo clamps to x in [0.02, 0.92],
    y in [0.10, 0.92] (/root/reference/app.mjs:362-367): exported cards must land on the canvas."""
    rng = np.random.default_rng(0)
    x = (rng.standard_normal((200, 2)) * [50.0, 0.001] + [1000.0, -3.0]).astype(np.float32)
    lab = rng.integers(0, 3, 200)
    b = bj.export_board(x, lab, np.zeros((3, 2)))
    xs = np.array([[b["meta"]["pos:p:%d" % i]["x"], b["meta"]["pos:p:%d" % i]["y"]] for i in range(200)])
    assert xs[:, 0].min() >= 0.02 and xs[:, 0].max() <= 0.92 and xs[:, 1].min() >= 0.10 and xs[:, 1].max() <= 0.92
    assert np.argmax(xs[:, 0]) == np.argmax(x[:, 0]) and np.argmin(xs[:, 1]) == np.argmin(x[:, 1])   # order kept
    back = bj.import_board(bj.loads(bj.dumps(b)))
    assert np.array_equal(back["points"], x)           # the raw coordinates still round-trip through `traits`
    one = bj.export_board(x[:1], lab[:1], np.zeros((3, 2)))
    assert one["meta"]["pos:p:0"] == pytest.approx({"x": 0.47, "y": 0.51})
