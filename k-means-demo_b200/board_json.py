"""Export / import in the reference's only on-disk format.

The demo's Export button writes `{cards, centroids, meta}` (/root/reference/app.mjs:263-267):
`cards` are `{id, title, traits, assignedTo, createdBy}` objects (`app.mjs:143-145`),
`centroids` are `{id, name, color, locked}` (`app.mjs:126-129`), `meta` is a flat map that holds
`iteration` (`app.mjs:288`), `mode`, and one `pos:<cardId> -> {x, y}` entry per placed card
(`app.mjs:158-159`).  Import replaces the board with the file's arrays and copies every meta
entry (`app.mjs:268-283`).

Mapping used here (ours — the reference has no numeric points):
  point i      -> card `{id: "p:<i>", title: "#<i>", traits: [repr(x_i0), repr(x_i1), ...],
                   assignedTo: "c:<label>" | null, createdBy: "kmeans_b200"}`;
                   for d == 2 also `meta["pos:p:<i>"] = {x, y}` — the reference's own place for a
                   card's position: a FRACTION of the drop zone, clamped by the demo to
                   x in [0.02, 0.92], y in [0.10, 0.92] (`app.mjs:362-367`), so the coordinates are
                   mapped affinely from the points' bounding box into that box (raw values stay in
                   `traits`, which is what import reads)
  cluster j    -> centroid `{id: "c:<j>", name: "Centroid <j+1>" (the default of app.mjs:128),
                   color, locked: false, coords: [...], count}`; `coords` and `count` are extra
                   keys the demo ignores
  iteration    -> `meta.iteration`; inertia / balance ride along as `meta.kmeans`
`repr()` of a Python float round-trips exactly, so export -> import reproduces the arrays bit
for bit.  A board exported by the demo itself (word traits) imports as labels + ids only.
"""
import json

import numpy as np

CREATED_BY = "kmeans_b200"


def _color(j):
    """A distinct colour per cluster: golden-angle steps around the hue circle (the demo picks
    "first unused of a fixed palette, else random", app.mjs:125; any CSS colour string works)."""
    return "hsl(%d, 70%%, 55%%)" % int((j * 137.50776405) % 360)


POS_BOX = ((0.02, 0.92), (0.10, 0.92))   # the demo's clamp of a card's position (app.mjs:366-367)


def canvas_positions(x):
    """d == 2 points -> fractions of the drop zone inside the demo's clamp box (bounding box of
    the points mapped affinely onto it; a degenerate axis sits in the middle)."""
    x = np.asarray(x, dtype=np.float64)
    out = np.empty_like(x)
    for t, (lo, hi) in enumerate(POS_BOX):
        a, b = float(x[:, t].min()), float(x[:, t].max())
        out[:, t] = (lo + hi) / 2 if b <= a else lo + (hi - lo) * (x[:, t] - a) / (b - a)
        out[:, t] = np.clip(out[:, t], lo, hi)
    return out


def export_board(points, labels, centroids, iteration=0, inertia=None, counts=None, mode="kmeans"):
    """Arrays -> the `{cards, centroids, meta}` dict the demo's Import accepts."""
    x = np.asarray(points)
    lab = np.asarray(labels, dtype=np.int64)
    c = np.asarray(centroids, dtype=np.float64)
    if x.ndim != 2 or c.ndim != 2 or x.shape[1] != c.shape[1] or lab.shape != (x.shape[0],):
        raise ValueError("need points [n][d], labels [n], centroids [k][d]")
    k = c.shape[0]
    if lab.size and (lab.min() < -1 or lab.max() >= k):
        raise ValueError("labels must lie in [-1, k)")
    if counts is None:
        counts = np.bincount(lab[lab >= 0], minlength=k)
    cards, meta = [], {"iteration": int(iteration), "mode": mode}
    pos = None
    if x.shape[1] == 2 and x.shape[0]:
        pos = canvas_positions(x)
    for i in range(x.shape[0]):
        cid = "p:%d" % i
        cards.append({"id": cid, "title": "#%d" % i, "traits": [repr(float(v)) for v in x[i]],
                      "assignedTo": None if lab[i] < 0 else "c:%d" % lab[i], "createdBy": CREATED_BY})
        if pos is not None:
            meta["pos:" + cid] = {"x": float(pos[i, 0]), "y": float(pos[i, 1])}
    cents = [{"id": "c:%d" % j, "name": "Centroid %d" % (j + 1), "color": _color(j), "locked": False,
              "coords": [float(v) for v in c[j]], "count": int(counts[j])} for j in range(k)]
    meta["kmeans"] = {"d": int(x.shape[1]), "dtype": str(x.dtype), "inertia": None if inertia is None else float(inertia)}
    return {"cards": cards, "centroids": cents, "meta": meta}


def import_board(board):
    """The inverse: `{cards, centroids, meta}` -> dict(points, labels, centroids, iteration, ids).

    `points` / `centroids` are None when the board carries no numbers (a board the demo itself
    exported: traits are words, centroids have no coordinates); labels are always recovered —
    the index of `assignedTo` in the centroid list, -1 for unassigned (or for an id that no
    longer exists: the demo un-assigns those cards when a centroid is removed, app.mjs:131-142)."""
    if not isinstance(board, dict) or not isinstance(board.get("cards"), list) or not isinstance(board.get("centroids"), list):
        raise ValueError("not a board: need {cards: [...], centroids: [...]}")
    meta = board.get("meta") if isinstance(board.get("meta"), dict) else {}
    cent_ids = [c.get("id") for c in board["centroids"]]
    index = {cid: j for j, cid in enumerate(cent_ids)}
    labels = np.array([index.get(card.get("assignedTo"), -1) for card in board["cards"]], dtype=np.int32)
    info = meta.get("kmeans") if isinstance(meta.get("kmeans"), dict) else {}
    dtype = np.dtype(info.get("dtype", "float32"))

    def numbers(seq):
        try:
            return [float(v) for v in seq]
        except (TypeError, ValueError):
            return None

    rows = [numbers(card.get("traits") or []) for card in board["cards"]]
    points = None
    if rows and all(r is not None for r in rows) and len({len(r) for r in rows}) == 1 and len(rows[0]) > 0:
        points = np.array(rows, dtype=np.float64).astype(dtype)
    crow = [c.get("coords") for c in board["centroids"]]
    centroids = np.array(crow, dtype=np.float64) if crow and all(isinstance(r, list) for r in crow) else None
    return {"points": points, "labels": labels, "centroids": centroids, "iteration": int(meta.get("iteration", 0) or 0),
            "card_ids": [card.get("id") for card in board["cards"]], "centroid_ids": cent_ids,
            "inertia": info.get("inertia")}


def dumps(board):
    """The demo writes `JSON.stringify(data, null, 2)` (app.mjs:265)."""
    return json.dumps(board, indent=2)


def loads(text):
    return json.loads(text)
