"""RowExpression IR mirror (core/trino-main/src/main/java/io/trino/sql/relational/):
CallExpression.java:28-55, SpecialForm.java:31-152, ConstantExpression.java:25-45,
InputReferenceExpression.java:25-48. Each node serialises to the plan language of DESIGN.md; anything
outside the GPU subset is rejected by the native plan compiler with NOT_SUPPORTED."""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Sequence

from .spi import BOOLEAN, Type


class RowExpression:
    type: Type

    def sexpr(self) -> str:
        raise NotImplementedError


@dataclass
class InputReferenceExpression(RowExpression):
    field: int
    type: Type

    def sexpr(self):
        return "(in %d %s)" % (self.field, self.type)


@dataclass
class ConstantExpression(RowExpression):
    value: object
    type: Type

    def sexpr(self):
        v = self.value
        if v is None:
            lit = "null"
        elif self.type.kind == 0:
            lit = "true" if v else "false"
        elif self.type.kind == 4:
            lit = "0x%016x" % struct.unpack("<Q", struct.pack("<d", float(v)))[0]
        else:
            lit = str(int(v))
        return "(const %s %s)" % (self.type, lit)


@dataclass
class CallExpression(RowExpression):
    """name: add subtract multiply divide modulus negate eq ne lt le gt ge not cast"""
    name: str
    type: Type
    arguments: Sequence[RowExpression]

    def sexpr(self):
        return "(call %s %s %s)" % (self.name, self.type, " ".join(a.sexpr() for a in self.arguments))


@dataclass
class SpecialForm(RowExpression):
    """form: AND OR BETWEEN IN IS_NULL IF COALESCE NULL_IF SWITCH WHEN (SpecialForm.Form, SpecialForm.java:137-152);
    SWITCH arguments = [value, WHEN(operand, result)..., optional else] as SwitchCodeGenerator.java:48-66 reads them"""
    form: str
    type: Type
    arguments: Sequence[RowExpression]

    def sexpr(self):
        args = " ".join(a.sexpr() for a in self.arguments)
        f = self.form.upper()
        if f == "AND":
            return "(and %s)" % args
        if f == "OR":
            return "(or %s)" % args
        if f == "BETWEEN":
            return "(between %s)" % args
        if f == "IN":
            return "(in-list %s)" % args
        if f == "IS_NULL":
            return "(is-null %s)" % args
        if f == "IF":
            return "(if %s %s)" % (self.type, args)
        if f == "COALESCE":
            return "(coalesce %s %s)" % (self.type, args)
        if f == "SWITCH":
            return "(switch %s %s)" % (self.type, args)
        # NULL_IF / WHEN as they are; anything else is forwarded verbatim so the native side reports NOT_SUPPORTED
        return "(%s %s)" % (f.lower().replace("_", "-"), args)


# small construction helpers (the Expressions / Signatures test utilities of the reference)
def field(i: int, t: Type) -> InputReferenceExpression:
    return InputReferenceExpression(i, t)


def constant(v, t: Type) -> ConstantExpression:
    return ConstantExpression(v, t)


def call(name: str, t: Type, *args: RowExpression) -> CallExpression:
    return CallExpression(name, t, list(args))


def compare(name: str, a: RowExpression, b: RowExpression) -> CallExpression:
    return CallExpression(name, BOOLEAN, [a, b])


def and_(*args: RowExpression) -> SpecialForm:
    return SpecialForm("AND", BOOLEAN, list(args))


def or_(*args: RowExpression) -> SpecialForm:
    return SpecialForm("OR", BOOLEAN, list(args))
