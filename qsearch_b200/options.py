"""A small Options container with the attribute / dict access pattern the reference's search code relies on
(qsearch/options.py:16-303): explicit values -> lazily evaluated smart defaults -> plain defaults.

Only what the hot path's callers need is here (no per-field pickling, no file I/O): ``options.target``,
``options.gateset``, ``options.solver`` ..., ``updated()``, ``set_defaults()``, ``set_smart_defaults()``,
``make_required()``, ``generate_cache()`` and ``in``.  The classes in this package also accept genuine
``qsearch.options.Options`` objects: they only ever use attribute access and ``in``.
"""


class Options:
    def __init__(self, defaults=None, smart_defaults=None, **values):
        d = self.__dict__
        d["_values"] = dict(values)
        d["_defaults"] = dict(defaults or {})
        d["_smart"] = dict(smart_defaults or {})
        d["_required"] = set()
        d["_cache"] = {}

    # -- lookup ----------------------------------------------------------------------------------------
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        d = self.__dict__
        if name in d["_values"]:
            return d["_values"][name]
        if name in d["_required"]:
            raise AttributeError("Missing required parameter {}".format(name))
        if name in d["_cache"]:
            return d["_cache"][name]
        if name in d["_smart"]:
            d["_required"].add(name)  # guards against a smart default that asks for itself
            try:
                value = d["_smart"][name](self)
            finally:
                d["_required"].discard(name)
            d["_cache"][name] = value
            return value
        if name in d["_defaults"]:
            return d["_defaults"][name]
        raise AttributeError("Could not find option {}".format(name))

    def __setattr__(self, name, value):
        self.__dict__["_values"][name] = value
        self.__dict__["_cache"].clear()

    __getitem__ = __getattr__
    __setitem__ = __setattr__

    def __contains__(self, name):
        d = self.__dict__
        return name in d["_values"] or name in d["_defaults"] or name in d["_smart"]

    # -- construction ------------------------------------------------------------------------------------
    def set_defaults(self, **kw):
        self.__dict__["_defaults"].update(kw)

    def set_smart_defaults(self, **kw):
        self.__dict__["_smart"].update(kw)
        self.__dict__["_cache"].clear()

    def make_required(self, *names):
        self.__dict__["_required"].update(names)
        for n in names:
            if n not in self.__dict__["_values"]:
                raise AttributeError("Missing required parameter {}".format(n))

    def generate_cache(self):
        for name in list(self.__dict__["_smart"]):
            if name not in self.__dict__["_values"]:
                try:
                    getattr(self, name)
                except AttributeError:
                    pass

    def copy(self):
        o = Options()
        for k in ("_values", "_defaults", "_smart"):
            o.__dict__[k] = dict(self.__dict__[k])
        o.__dict__["_required"] = set(self.__dict__["_required"])
        return o

    def updated(self, other=None, **values):
        o = self.copy()
        o.update(other, **values)
        return o

    def update(self, other=None, **values):
        d = self.__dict__
        if other is not None:
            od = other.__dict__
            if "_values" in od and "_smart" in od:  # one of ours
                d["_defaults"].update(od["_defaults"])
                d["_smart"].update(od["_smart"])
                d["_values"].update(od["_values"])
            else:  # a qsearch.options.Options
                d["_defaults"].update(getattr(other, "defaults", {}))
                d["_smart"].update(getattr(other, "smart_defaults", {}))
                d["_values"].update({k: v for k, v in od.items() if k not in ("defaults", "smart_defaults", "required", "cache")})
        d["_values"].update(values)
        d["_cache"].clear()

    def __getstate__(self):
        return {k: self.__dict__[k] for k in ("_values", "_defaults", "_smart", "_required")}

    def __setstate__(self, st):
        self.__dict__.update(st)
        self.__dict__["_cache"] = {}
