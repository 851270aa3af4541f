"""Kubernetes resource.Quantity arithmetic, the subset needed to sum
container requests/limits into a PodGroup's minResources
(controllers/paddlejob_helper.go:520-548 in the reference) and to count GPUs.

Values are kept as exact fractions; formatting picks the suffix family of the
first operand the way resource.Quantity does (binary Ki/Mi/Gi, decimal k/M/G,
milli `m`)."""
from __future__ import annotations

from fractions import Fraction
from typing import Union

_BIN = {"Ki": 2 ** 10, "Mi": 2 ** 20, "Gi": 2 ** 30, "Ti": 2 ** 40, "Pi": 2 ** 50, "Ei": 2 ** 60}
_DEC = {"n": Fraction(1, 10 ** 9), "u": Fraction(1, 10 ** 6), "m": Fraction(1, 1000), "": 1, "k": 10 ** 3,
        "M": 10 ** 6, "G": 10 ** 9, "T": 10 ** 12, "P": 10 ** 15, "E": 10 ** 18}


class Quantity:
    __slots__ = ("value", "family")

    def __init__(self, value: Fraction, family: str = "dec"):
        self.value = Fraction(value)
        self.family = family  # "bin" | "dec"

    @classmethod
    def parse(cls, s: Union[str, int, float, "Quantity"]) -> "Quantity":
        if isinstance(s, Quantity):
            return cls(s.value, s.family)
        if isinstance(s, (int, float)):
            return cls(Fraction(str(s)), "dec")
        t = str(s).strip()
        if not t:
            raise ValueError("empty quantity")
        for suf, mul in _BIN.items():
            if t.endswith(suf):
                return cls(Fraction(t[: -len(suf)]) * mul, "bin")
        # scientific notation: 1e3 / 1E3
        low = t.lower()
        if "e" in low and not t.endswith("E"):
            mant, exp = low.split("e", 1)
            return cls(Fraction(mant) * Fraction(10) ** int(exp), "dec")
        suf = t[-1] if t[-1] in "numkMGTPE" else ""
        num = t[: -1] if suf else t
        return cls(Fraction(num) * _DEC[suf], "dec")

    def __add__(self, other: "Quantity") -> "Quantity":
        return Quantity(self.value + Quantity.parse(other).value, self.family)

    def __eq__(self, other) -> bool:
        return self.value == Quantity.parse(other).value

    def __le__(self, other) -> bool:
        return self.value <= Quantity.parse(other).value

    def __lt__(self, other) -> bool:
        return self.value < Quantity.parse(other).value

    def __int__(self) -> int:
        return int(self.value)

    def __float__(self) -> float:
        return float(self.value)

    def __repr__(self) -> str:
        return f"Quantity({self})"

    def __str__(self) -> str:
        v = self.value
        if v == 0:
            return "0"
        if self.family == "bin":
            for suf in ("Ei", "Pi", "Ti", "Gi", "Mi", "Ki"):
                if v % _BIN[suf] == 0:
                    return f"{v // _BIN[suf]}{suf}"
            if v.denominator == 1:
                return str(v.numerator)
        if v.denominator == 1:
            n = v.numerator
            for suf in ("E", "P", "T", "G", "M", "k"):
                if n % _DEC[suf] == 0:
                    return f"{n // _DEC[suf]}{suf}"
            return str(n)
        milli = v * 1000
        if milli.denominator == 1:
            return f"{milli.numerator}m"
        return str(float(v))


def add_resource_lists(total: dict, res: dict) -> None:
    """total[name] += res[name] for every name (addRes in the reference)."""
    for name, q in (res or {}).items():
        if name in total:
            total[name] = Quantity.parse(total[name]) + Quantity.parse(q)
        else:
            total[name] = Quantity.parse(q)


def resource_list_to_strings(rl: dict) -> dict:
    return {k: str(Quantity.parse(v)) for k, v in rl.items()}
