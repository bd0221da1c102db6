"""flatten / inflate: path grammar, order and container entries (reference flatten.py:18-224), with the shapes the
fast paths of this build's flatten see: wide dicts of leaves, keys that need escaping, mixed key types."""
from collections import OrderedDict

import torch

import tsb200_loader

tsb200_loader.load()
from torchsnapshot_b200.flatten import flatten, inflate  # noqa: E402


def test_wide_dict_of_leaves_keeps_order_and_objects():
    d = {f"layer{i}.weight": torch.zeros(1) for i in range(200)}
    manifest, flat = flatten(d, "model")
    assert list(manifest) == ["model"] and manifest["model"].keys == list(d)
    assert list(flat) == [f"model/{k}" for k in d]
    assert all(flat[f"model/{k}"] is v for k, v in d.items())


def test_escaping_in_wide_and_narrow_dicts():
    wide = {f"a/{i}%": i for i in range(40)}
    _m, flat = flatten({"w": wide, "n": {"x/y": 1}}, "p/q")
    assert list(flat)[:2] == ["p%2Fq/w/a%2F0%25", "p%2Fq/w/a%2F1%25"]
    assert "p%2Fq/n/x%2Fy" in flat
    assert inflate(_m, flat, "p/q") == {"w": wide, "n": {"x/y": 1}}


def test_preorder_with_containers_between_leaves():
    obj = OrderedDict([("a", 1), ("b", [10, {"c": 2}]), ("d", {str(i): i for i in range(12)}), ("e", 3)])
    manifest, flat = flatten(obj, "k")
    assert list(flat) == ["k/a", "k/b/0", "k/b/1/c"] + [f"k/d/{i}" for i in range(12)] + ["k/e"]
    assert list(manifest) == ["k", "k/b", "k/b/1", "k/d"]
    assert type(manifest["k"]).__name__ == "OrderedDictEntry" and type(manifest["k/d"]).__name__ == "DictEntry"
    assert inflate(manifest, flat, "k") == obj


def test_int_keys_and_colliding_keys():
    ints = {i: i * i for i in range(20)}
    manifest, flat = flatten({"s": ints, "bad": {1: "a", "1": "b"}, "odd": {(1, 2): 3}}, "k")
    assert list(flat)[:3] == ["k/s/0", "k/s/1", "k/s/2"]
    assert flat["k/bad"] == {1: "a", "1": "b"} and flat["k/odd"] == {(1, 2): 3}  # opaque leaves
    out = inflate(manifest, flat, "k")
    assert out["s"] == ints and list(out["s"])[0] == 0  # ints stay ints


def test_empty_containers_and_scalars():
    manifest, flat = flatten({"e": {}, "l": [], "x": None}, "k")
    assert list(manifest) == ["k", "k/e", "k/l"] and list(flat) == ["k/x"]
    assert inflate(manifest, flat, "k") == {"e": {}, "l": [], "x": None}
    manifest, flat = flatten(5, "k")
    assert manifest == {} and flat == {"k": 5}


def test_wide_dicts_get_their_paths_back_and_follow_changes():
    d = {f"w{i}": i for i in range(3000)}
    _m, first = flatten(d, "big")
    _m, again = flatten(d, "big")
    assert list(again) == list(first) and all(a is b for a, b in zip(again, first))  # the same string objects
    again.pop("big/w0")  # the caller's dict is its own
    assert "big/w0" in flatten(d, "big")[1]
    d["extra/1"] = -1
    del d["w5"]
    _m, changed = flatten(d, "big")
    assert "big/w5" not in changed and changed["big/extra%2F1"] == -1 and len(changed) == 3000
    assert list(changed)[:3] == ["big/w0", "big/w1", "big/w2"]
    _m, other = flatten({f"v{i}": i for i in range(3000)}, "big")  # another dict under the same prefix
    assert list(other)[0] == "big/v0"


def _reference_walk(obj, path, manifest, flat):
    """The grammar restated recursively (reference flatten.py:18-120): containers first, children in order."""
    if type(obj) is list:
        manifest[path] = "list"
        for i, v in enumerate(obj):
            _reference_walk(v, f"{path}/{i}", manifest, flat)
    elif type(obj) in (dict, OrderedDict) and all(isinstance(k, (str, int)) for k in obj) \
            and len({str(k) for k in obj}) == len(obj):
        manifest[path] = ("OrderedDict" if type(obj) is OrderedDict else "dict", list(obj.keys()))
        for k, v in obj.items():
            enc = str(k).replace("%", "%25").replace("/", "%2F")
            _reference_walk(v, f"{path}/{enc}", manifest, flat)
    else:
        flat[path] = obj


def test_random_trees_against_the_recursive_restatement():
    import random

    rnd = random.Random(20)

    def key():
        kind = rnd.random()
        if kind < 0.15:
            return rnd.randrange(50)
        alphabet = "abc/%._0"
        return "".join(rnd.choice(alphabet) for _ in range(rnd.randrange(1, 6)))

    def tree(depth):
        kind = rnd.random()
        if depth == 0 or kind < 0.35:
            return rnd.choice([1, 2.5, "s", None, (1, 2), torch.zeros(1)])
        if kind < 0.55:
            return [tree(depth - 1) for _ in range(rnd.randrange(0, 4))]
        n = rnd.choice([0, 1, 3, 12, 40])
        d = OrderedDict() if rnd.random() < 0.3 else {}
        for _ in range(n):
            d[key()] = tree(depth - 1) if rnd.random() < 0.4 else rnd.randrange(100)
        return d

    for _ in range(300):
        obj = tree(4)
        prefix = rnd.choice(["k", "a/b", "p%q"])
        manifest, flat = flatten(obj, prefix)
        want_m, want_f = {}, {}
        _reference_walk(obj, prefix.replace("%", "%25").replace("/", "%2F"), want_m, want_f)
        assert list(flat) == list(want_f) and all(flat[p] is want_f[p] for p in flat)
        assert list(manifest) == list(want_m)
        for p, e in manifest.items():
            assert (e.type == "list") if want_m[p] == "list" else (e.type, e.keys) == want_m[p]
        back = inflate(manifest, flat, prefix)

        def same(a, b):
            if type(a) is not type(b):
                return False
            if isinstance(a, dict):
                return list(a) == list(b) and all(same(a[k], b[k]) for k in a)
            if isinstance(a, list):
                return len(a) == len(b) and all(same(x, y) for x, y in zip(a, b))
            return a is b or a == b

        assert same(back, obj)
