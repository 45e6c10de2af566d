"""CPU restatement of the reference's dashboard metrics — TEST INFRASTRUCTURE, NOT PRODUCT.

This is the ONE piece of arithmetic the reference actually contains, restated line by line
from /root/reference/app.mjs:
    normTokens            app.mjs:436-443
    titleCase             app.mjs:444
    tokensForCard         app.mjs:445-449
    traitCountsFor        app.mjs:450-461
    cohesionFor           app.mjs:462-475
    suggestionFromCounts  app.mjs:476-480
    snapshotMetrics       app.mjs:481-496
and pinned against the reference's own seed data (JESSICA app.mjs:188, populateTestData
app.mjs:204-216) with the known answers of SURVEY.md Appendix B (tests/test_cluster_metrics.py).
It pins row a6/f2 of the coverage table (counts, balance, cohesion), not the Lloyd path.
"""
import math
import re

# app.mjs:188 and app.mjs:204-216 (id, title, traitA, traitB)
SEED_CARDS = [
    ("seed:jessica", "Jessica", "Fresh", "Sorbet"),
    ("seed:t1", "Nguyen", "Sweet", "Creamy"),
    ("seed:t2", "Patel", "Fresh", "Sorbet"),
    ("seed:t3", "Garcia", "Chocolatey", "Crunchy"),
    ("seed:t4", "Rossi", "Milky", "Silky"),
    ("seed:t5", "Kim", "Nutty", "Creamy"),
    ("seed:t6", "Smith", "Fruity", "Swirled"),
    ("seed:t7", "Ahmed", "Bitter", "Rich"),
    ("seed:t8", "Lopez", "Sweet", "Colorful"),
    ("seed:t9", "Chen", "Rich", "Spicy"),
    ("seed:t10", "Nils", "Espresso", "Hot"),
    ("seed:t11", "sally", "Vegan", "Not Sweet"),
]


def seed_cards():
    return [dict(id=i, title=t, traits=[a, b], assignedTo=None) for i, t, a, b in SEED_CARDS]


_SPLIT = re.compile(r"[/,&•+]|(?:\s+and\s+)|\||,", re.IGNORECASE)


def norm_tokens(s):
    """app.mjs:436-443 — split on / , & • + | and ' and ', trim, drop empties, lower-case."""
    if not s:
        return []
    return [x.strip().lower() for x in _SPLIT.split(str(s)) if x.strip()]


def title_case(s):
    """app.mjs:444 — /\\w\\S*/g, first char upper-cased, rest untouched."""
    return re.sub(r"\w\S*", lambda m: m.group(0)[0].upper() + m.group(0)[1:], s)


def tokens_for_card(card):
    """app.mjs:445-449 — both traits, de-duplicated within the card (insertion order kept)."""
    traits = (card or {}).get("traits") or [None, None]
    out = []
    for t in norm_tokens(traits[0] if len(traits) > 0 else None) + norm_tokens(traits[1] if len(traits) > 1 else None):
        if t not in out:
            out.append(t)
    return out


def trait_counts_for(cards):
    """app.mjs:450-461 — token -> {label, count}."""
    m = {}
    for c in cards:
        for t in tokens_for_card(c):
            e = m.get(t) or dict(label=title_case(t), count=0)
            e["count"] += 1
            m[t] = e
    return m


def cohesion_for(cards):
    """app.mjs:462-475 — share of cards that share a token with another card; 1 when n <= 1."""
    n = len(cards)
    if n <= 1:
        return 1
    sets = [set(tokens_for_card(c)) for c in cards]
    share = 0
    for i in range(n):
        if any(i != j and sets[i] & sets[j] for j in range(n)):
            share += 1
    return share / n


def suggestion_from_counts(m):
    """app.mjs:476-480 — top two by (count desc, label asc)."""
    arr = sorted(m.values(), key=lambda e: (-e["count"], e["label"]))
    if not arr:
        return None
    return f"{arr[0]['label']} + {arr[1]['label']}" if len(arr) > 1 else arr[0]["label"]


def snapshot_metrics(cards, centroids):
    """app.mjs:481-496 — counts, cohesion, balance{max,min,gap,ratio}, avgCohesion."""
    counts, coh = {}, {}
    for c in centroids:
        cs = [x for x in cards if x.get("assignedTo") == c["id"]]
        counts[c["id"]] = len(cs)
        coh[c["id"]] = cohesion_for(cs)
    vals = list(counts.values())
    mx = max(vals) if vals else 0
    mn = min(vals) if vals else 0
    ratio = (mx / mn) if mn else (math.inf if mx else 1)
    avg = (sum(coh.values()) / len(coh)) if coh else 1
    return dict(counts=counts, cohesion=coh, balance=dict(max=mx, min=mn, gap=mx - mn, ratio=ratio), avgCohesion=avg)
