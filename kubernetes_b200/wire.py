"""Python writer for the C-ABI's wire records (include/ksb_wire.hpp), from the same dict shape
as the JSON interchange (csrc/host/model.hpp). Lets tests drive libksched_b200.so directly through
ctypes, the way the Go shim would through cgo."""
import struct


class W:
    def __init__(self):
        self.b = bytearray()

    def u8(self, v):
        self.b += struct.pack("<B", 1 if v else 0)

    def u32(self, v):
        self.b += struct.pack("<I", v)

    def i32(self, v):
        self.b += struct.pack("<i", v)

    def i64(self, v):
        self.b += struct.pack("<q", int(v))

    def s(self, v):
        d = (v or "").encode()
        self.u32(len(d))
        self.b += d

    def map(self, m):
        m = m or {}
        self.u32(len(m))
        for k in sorted(m):
            self.s(k)
            self.s(m[k])

    def reslist(self, m):
        m = m or {}
        self.u32(len(m))
        for k in sorted(m):
            self.s(k)
            self.i64(m[k])

    def strs(self, v):
        v = v or []
        self.u32(len(v))
        for x in v:
            self.s(x)

    def reqs(self, v):
        v = v or []
        self.u32(len(v))
        for r in v:
            self.s(r.get("key", ""))
            self.s(r.get("op", ""))
            self.strs(r.get("values"))


def node_record(n):
    w = W()
    w.s(n["name"])
    w.u8(n.get("labels") is not None)
    w.map(n.get("labels"))
    w.reslist(n.get("allocatable"))
    taints = n.get("taints") or []
    w.u32(len(taints))
    for t in taints:
        w.s(t.get("key"))
        w.s(t.get("value"))
        w.s(t.get("effect"))
    conds = n.get("conditions") or []
    w.u32(len(conds))
    for c in conds:
        w.s(c["type"])
        w.s(c["status"])
    w.u8(n.get("unschedulable", False))
    w.u8(n.get("avoidInvalid", False))
    avoid = n.get("avoid") or []
    w.u32(len(avoid))
    for a in avoid:
        w.s(a["kind"])
        w.s(a["uid"])
    ext = n.get("extendedResources") or {}
    w.u32(len(ext))
    for name in sorted(ext):
        w.s(name)
        w.u32(len(ext[name]))
        for d in ext[name]:
            w.s(d["id"])
            w.s(d.get("health", ""))
            w.map(d.get("attributes"))
    images = n.get("images") or []
    w.u32(len(images))
    for im in images:
        w.strs(im.get("names"))
        w.i64(im.get("sizeBytes", 0))
    return bytes(w.b)


def _container(w, c):
    w.reslist(c.get("requests"))
    w.reslist(c.get("limits"))
    ports = c.get("ports") or []
    w.u32(len(ports))
    for p in ports:
        w.i32(p.get("hostPort", 0))
        w.s(p.get("protocol", ""))
        w.s(p.get("hostIP", ""))


def _term(w, t):
    sel = t.get("selector")
    w.u8(sel is not None)
    w.map((sel or {}).get("matchLabels"))
    w.reqs((sel or {}).get("matchExpressions"))
    w.strs(t.get("namespaces"))
    w.s(t.get("topologyKey", ""))


def _pod_affinity(w, a):
    w.u8(a is not None)
    a = a or {}
    req = a.get("required") or []
    w.u32(len(req))
    for t in req:
        _term(w, t)
    pref = a.get("preferred") or []
    w.u32(len(pref))
    for t in pref:
        w.i32(t.get("weight", 0))
        _term(w, t.get("term") or {})


def pod_record(p):
    w = W()
    w.s(p.get("namespace", ""))
    w.s(p.get("name", ""))
    w.map(p.get("labels"))
    w.u8(p.get("deleting", False))
    w.s(p.get("nodeName", ""))
    ctl = p.get("controller")
    w.u8(ctl is not None)
    w.s((ctl or {}).get("kind", ""))
    w.s((ctl or {}).get("uid", ""))
    for key in ("containers", "initContainers"):
        cs = p.get(key) or []
        w.u32(len(cs))
        for c in cs:
            _container(w, c)
    w.map(p.get("nodeSelector"))
    aff = p.get("affinity")
    w.u8(aff is not None)
    aff = aff or {}
    na = aff.get("node")
    w.u8(na is not None)
    na = na or {}
    req = na.get("required")
    w.u8(req is not None)
    req = req or []
    w.u32(len(req))
    for t in req:
        w.reqs(t.get("matchExpressions"))
    pref = na.get("preferred") or []
    w.u32(len(pref))
    for t in pref:
        w.i32(t.get("weight", 0))
        w.reqs(t.get("matchExpressions"))
    _pod_affinity(w, aff.get("podAffinity"))
    _pod_affinity(w, aff.get("podAntiAffinity"))
    tols = p.get("tolerations") or []
    w.u32(len(tols))
    for t in tols:
        w.s(t.get("key", ""))
        w.s(t.get("operator", ""))
        w.s(t.get("value", ""))
        w.s(t.get("effect", ""))
    ext = p.get("extendedResources") or []
    w.u32(len(ext))
    for e in ext:
        w.s(e.get("name", ""))
        w.reslist(e.get("limits"))
        w.reqs(e.get("required"))
        w.strs(e.get("assigned"))
    w.i32(p.get("volumes", 0))
    w.i32(p.get("priority", 0))
    w.strs([c.get("image", "") for c in (p.get("containers") or [])])
    return bytes(w.b)


def owner_record(o, label_selector_kind):
    w = W()
    w.s(o.get("namespace", ""))
    w.s(o.get("name", ""))
    if label_selector_kind:
        w.map(None)
        sel = o.get("selector")
        w.u8(sel is not None)
        w.map((sel or {}).get("matchLabels"))
        w.reqs((sel or {}).get("matchExpressions"))
    else:
        w.map(o.get("selector"))
        w.u8(False)
        w.map(None)
        w.reqs(None)
    return bytes(w.b)
