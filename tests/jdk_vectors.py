"""Known answers PUBLISHED for the primitives the reference delegates to the JDK (java.base, JDK 11, pom.xml:13-15):
the reference's own tests pin only `1.0 + 2.0` and `10.0 * null` at this boundary, so the oracle and
the kernels are additionally held to the examples and rules printed in the Java Language Specification (SE 11) and
the java.lang Javadoc.  Call sites in the reference: DREM / DDIV `M/evaluator/BytecodeCompiler.kt:328-332`,
`Double.compare` `:300` and `M/evaluator/Interpreter.kt:104-107`, boxed `equals` `Interpreter.kt:102-103`,
`Math.min/max` `M/evaluator/Accumulators.kt:62,80`.

Each entry: (Function name, a, b, expected) with expected a float (NaN compares as NaN, zeros by sign) or a bool.
"""
nan, inf = float("nan"), float("inf")

ARITHMETIC = [
    # JLS 15.17.3, the printed examples: "5.0%3.0 produces 2.0, 5.0%(-3.0) produces 2.0, (-5.0)%3.0 produces -2.0,
    # (-5.0)%(-3.0) produces -2.0"
    ("MOD", 5.0, 3.0, 2.0), ("MOD", 5.0, -3.0, 2.0), ("MOD", -5.0, 3.0, -2.0), ("MOD", -5.0, -3.0, -2.0),
    # JLS 15.17.3 rules: NaN operand -> NaN; infinite dividend or zero divisor -> NaN; finite % infinite -> dividend;
    # zero % finite -> dividend (with its sign)
    ("MOD", nan, 3.0, nan), ("MOD", 5.0, nan, nan), ("MOD", inf, 3.0, nan), ("MOD", -inf, 3.0, nan), ("MOD", 5.0, 0.0, nan),
    ("MOD", 5.0, -0.0, nan), ("MOD", 5.0, inf, 5.0), ("MOD", -5.0, -inf, -5.0), ("MOD", 0.0, 3.0, 0.0), ("MOD", -0.0, 3.0, -0.0),
    # the sign of the result equals the sign of the dividend, also when the result is zero
    ("MOD", -6.0, 3.0, -0.0), ("MOD", 6.0, -3.0, 0.0),
    # JLS 15.17.2: nonzero / zero -> signed infinity; zero / zero and inf / inf -> NaN; finite / inf -> signed zero
    ("DIV", 1.0, 0.0, inf), ("DIV", -1.0, 0.0, -inf), ("DIV", 1.0, -0.0, -inf), ("DIV", 0.0, 0.0, nan), ("DIV", inf, inf, nan),
    ("DIV", 1.0, inf, 0.0), ("DIV", -1.0, inf, -0.0), ("DIV", inf, 2.0, inf),
    # JLS 15.17.1 / 15.18.2: inf * 0 -> NaN; inf - inf -> NaN; opposite zeros add to +0.0; equal zeros keep their sign
    ("MUL", inf, 0.0, nan), ("MUL", -2.0, inf, -inf), ("SUB", inf, inf, nan), ("ADD", inf, -inf, nan), ("ADD", 0.0, -0.0, 0.0),
    ("ADD", -0.0, -0.0, -0.0), ("SUB", 1.0, 1.0, 0.0),
    # round to nearest even, no fused multiply-add (JLS 15.17 / 4.2.4): the classic
    ("ADD", 0.1, 0.2, 0.30000000000000004), ("MUL", 0.1, 3.0, 0.30000000000000004), ("SUB", 0.3, 0.1, 0.19999999999999998),
]

COMPARISONS = [
    # Double.compare Javadoc: "Double.NaN is considered by this method to be equal to itself and greater than all other
    # double values (including Double.POSITIVE_INFINITY). 0.0d is considered by this method to be greater than -0.0d."
    ("CMP_GT", nan, inf, True), ("CMP_GE", nan, inf, True), ("CMP_LT", inf, nan, True), ("CMP_LE", nan, nan, True),
    ("CMP_GE", nan, nan, True), ("CMP_LT", nan, nan, False), ("CMP_GT", nan, nan, False), ("CMP_GT", 0.0, -0.0, True),
    ("CMP_LT", -0.0, 0.0, True), ("CMP_LE", 0.0, -0.0, False), ("CMP_GE", -0.0, 0.0, False), ("CMP_LT", -inf, -1.7976931348623157e308, True),
    ("CMP_LT", 1.0, 2.0, True), ("CMP_GE", 2.0, 2.0, True),
    # Double.equals Javadoc: NaN equals NaN, +0.0 does not equal -0.0 (Kotlin == on boxed values, Interpreter.kt:102-103)
    ("CMP_EQ", nan, nan, True), ("CMP_NE", nan, nan, False), ("CMP_EQ", 0.0, -0.0, False), ("CMP_NE", 0.0, -0.0, True),
    ("CMP_EQ", 1.5, 1.5, True), ("CMP_NE", 1.5, 2.5, True),
]

# Math.min / Math.max Javadoc: "If either value is NaN, then the result is NaN. Unlike the numerical comparison
# operators, this method considers negative zero to be strictly smaller than positive zero."
# (column values, expected MIN, expected MAX)
MIN_MAX = [
    ([0.0, -0.0], -0.0, 0.0), ([-0.0, 0.0], -0.0, 0.0), ([1.0, nan, 2.0], nan, nan), ([nan], nan, nan),
    ([-inf, inf, 3.0], -inf, inf), ([5.0, 5.0], 5.0, 5.0),
]


def same(a, b):
    import math
    if isinstance(b, bool) or isinstance(a, bool):
        return a is b
    if math.isnan(b):
        return isinstance(a, float) and math.isnan(a)
    return a == b and math.copysign(1.0, a) == math.copysign(1.0, b)
