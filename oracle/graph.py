"""CPU oracle of one pass over a small inference graph, in numpy.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates the walk of
``engine/src/main/java/io/seldon/engine/predictors/PredictiveUnitBean.java:113-197`` (getOutputAsync: transformInput,
route, children, aggregate, transformOutput) with the meta rules of ``mergeMeta`` (L354-372: tags of the earlier
message(s), then the latest message's tags over them; metrics cleared from the message and collected per unit by
``addMetrics``, L96-110), the routing check of L345-351 and the routing / requestPath bookkeeping of L123 / L159.
The engine cannot be compiled here (no JVM); its only known answers on this path are the routing values of
``RandomABTestUnitInternalTest.java:59-74`` (checked in tests/test_oracle_graph.py through oracle/abtest.py).

A node is a dict {"name", "kind", "impl", "children", "image"}; impl objects are the numpy oracles of the components.
"""


def _meta(impl, name, metrics):
    tags = {}
    if hasattr(impl, "tags"):
        tags = impl.tags() or {}
    if hasattr(impl, "metrics"):
        m = impl.metrics()
        if m:
            metrics.setdefault(name, []).extend(m)
    return tags


def walk(node, X, tags_in, routing, path, metrics):
    path[node["name"]] = node.get("image", "")
    impl, kind = node["impl"], node["kind"]
    cur, tags = X, dict(tags_in)
    if kind in ("MODEL", "TRANSFORMER"):
        cur = impl.predict(X, None) if kind == "MODEL" else impl.transform_input(X, None)
        tags.update(_meta(impl, node["name"], metrics))
    children = node.get("children", [])
    if not children:
        return cur, tags
    branch = -1
    if kind == "ROUTER":
        branch = int(impl.route(cur, None))
        if branch < -1 or branch >= len(children):
            raise ValueError("Invalid branch index")
    routing[node["name"]] = branch
    selected = children if branch == -1 else [children[branch]]
    outs, child_tags = [], []
    for ch in selected:
        o, t = walk(ch, cur, tags, routing, path, metrics)
        outs.append(o)
        child_tags.append(t)
    if kind == "COMBINER":
        agg = impl.aggregate(outs, None)
        latest = _meta(impl, node["name"], metrics)
    else:
        assert len(outs) == 1
        agg, latest = outs[0], {}
    merged = {}
    for t in child_tags:
        merged.update(t)
    merged.update(latest)
    if kind == "OUTPUT_TRANSFORMER":
        agg = impl.transform_output(agg, None)
        merged.update(_meta(impl, node["name"], metrics))
    return agg, merged


def predict(root, X, puid=""):
    routing, path, metrics = {}, {}, {}
    out, tags = walk(root, X, {}, routing, path, metrics)
    return out, {"puid": puid, "tags": tags, "routing": routing, "requestPath": path, "metrics": metrics}
