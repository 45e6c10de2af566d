"""CPU: the reference's only arithmetic (dashboard metrics, app.mjs:436-496) restated in
oracle/cluster_metrics.py, checked against SURVEY.md Appendix B — values obtained by
hand-executing the reference on its own seed cards (app.mjs:188, 204-216).  These are the
only reference-derived known answers that exist; they pin the counts/balance/cohesion row
(a6/f2), and the same balance formulas are what km_get_balance implements on the GPU."""
import math

import oracle
from oracle import cluster_metrics as cm


def _assign(names_by_cluster):
    cards = cm.seed_cards()
    cents = [dict(id=f"c:{i}", name=f"Centroid {i + 1}") for i in range(len(names_by_cluster))]
    for i, names in enumerate(names_by_cluster):
        for c in cards:
            if c["title"] in names:
                c["assignedTo"] = cents[i]["id"]
    return cards, cents


def test_tokeniser_edges():
    assert cm.norm_tokens("Sweet and Salty / Nutty & rich|x+y•z") == ["sweet", "salty", "nutty", "rich", "x", "y", "z"]
    assert cm.norm_tokens("Not Sweet") == ["not sweet"]          # no whitespace split: does NOT match "sweet"
    assert cm.norm_tokens(None) == [] and cm.norm_tokens("") == []
    assert cm.title_case("not sweet") == "Not Sweet"
    assert len({t for c in cm.seed_cards() for t in cm.tokens_for_card(c)}) == 19


def test_all_twelve_in_one_cluster():
    cards = cm.seed_cards()
    assert cm.cohesion_for(cards) == 7 / 12 == 0.5833333333333334
    sharers = {c["title"] for i, c in enumerate(cards)
               if any(i != j and set(cm.tokens_for_card(c)) & set(cm.tokens_for_card(o)) for j, o in enumerate(cards))}
    assert sharers == {"Jessica", "Nguyen", "Patel", "Kim", "Ahmed", "Lopez", "Chen"}
    counts = cm.trait_counts_for(cards)
    top3 = sorted(counts.values(), key=lambda e: (-e["count"], e["label"]))[:3]
    assert [(e["label"], e["count"]) for e in top3] == [("Creamy", 2), ("Fresh", 2), ("Rich", 2)]
    assert cm.suggestion_from_counts(counts) == "Creamy + Fresh"


def test_three_clusters_of_four():
    cards, cents = _assign([{"Jessica", "Patel", "Smith", "sally"}, {"Nguyen", "Kim", "Lopez", "Rossi"},
                            {"Garcia", "Ahmed", "Chen", "Nils"}])
    m = cm.snapshot_metrics(cards, cents)
    assert list(m["counts"].values()) == [4, 4, 4]
    assert m["balance"] == dict(max=4, min=4, gap=0, ratio=1)
    assert list(m["cohesion"].values()) == [0.5, 0.75, 0.5]
    assert m["avgCohesion"] == 0.5833333333333334
    sugg = [cm.suggestion_from_counts(cm.trait_counts_for([c for c in cards if c["assignedTo"] == z["id"]])) for z in cents]
    assert sugg == ["Fresh + Sorbet", "Creamy + Sweet", "Bitter + Rich"] or sugg == ["Fresh + Sorbet", "Creamy + Sweet", "Rich + Bitter"]


def test_round_robin_has_no_cohesion():
    cards = cm.seed_cards()
    cents = [dict(id=f"c:{i}") for i in range(3)]
    for i, c in enumerate(cards):
        c["assignedTo"] = cents[i % 3]["id"]
    m = cm.snapshot_metrics(cards, cents)
    assert list(m["cohesion"].values()) == [0, 0, 0] and m["avgCohesion"] == 0


def test_edges_and_balance_match_the_lloyd_oracle():
    assert cm.cohesion_for([]) == 1 and cm.cohesion_for([cm.seed_cards()[0]]) == 1     # app.mjs:463
    cards, cents = _assign([{"Jessica", "Patel"}, set()])
    b = cm.snapshot_metrics(cards, cents)["balance"]
    assert b["ratio"] == math.inf and b["gap"] == 2                                   # app.mjs:493
    assert cm.snapshot_metrics([], [dict(id="a"), dict(id="b")])["balance"]["ratio"] == 1
    assert cm.snapshot_metrics([], [])["avgCohesion"] == 1
    # the numeric oracle's balance (and km_get_balance) use the same formulas
    for counts in ([2, 0], [0, 0], [4, 4, 4], [6, 3]):
        ids = [dict(id=str(i)) for i in range(len(counts))]
        cs = [dict(traits=["a", "b"], assignedTo=str(i)) for i, n in enumerate(counts) for _ in range(n)]
        assert cm.snapshot_metrics(cs, ids)["balance"] == oracle.balance(counts)
