"""Go encoding/json output rules restated by the oracle, checked two independent ways.

No reference test asserts the bytes of State.Bytes() (SURVEY.md 8c "not pinned"), so these
vectors are HAND-DERIVED from the published encoding/json rules, and the bulk of the format
(key order, indentation, separators, empty containers) is cross-checked against CPython's
json module, an independent implementation.
"""
import json
import random
import struct

import pytest

from oracle.oracle import OracleError


def py_go_bytes(obj) -> bytes:
    """What Go's MarshalIndent(v, "", "\\t") gives for string/list/dict/None/bool-only trees."""
    s = json.dumps(obj, indent="\t", sort_keys=True, ensure_ascii=False, separators=(",", ": "))
    out = []
    # post-process string escapes that differ: Go escapes <, >, &, U+2028/9 and DEL stays raw;
    # CPython writes \b and \f short forms where Go (<= 1.21) writes \u0008 and \u000c.
    s = (s.replace("<", "\\u003c").replace(">", "\\u003e").replace("&", "\\u0026")
          .replace("\u2028", "\\u2028").replace("\u2029", "\\u2029")
          .replace("\\b", "\\u0008").replace("\\f", "\\u000c"))
    return s.encode("utf-8")


def rand_tree(rng, depth=0):
    r = rng.random()
    alphabet = "abcXYZ019_-.${}/:~ é√\u2028<>&'\t\n"
    def rs():
        return "".join(rng.choice(alphabet) for _ in range(rng.randrange(0, 12)))
    if depth > 3 or r < 0.45:
        return rng.choice([rs(), rs(), rs(), None, True, False])
    if r < 0.6:
        return [rand_tree(rng, depth + 1) for _ in range(rng.randrange(0, 4))]
    return {rs(): rand_tree(rng, depth + 1) for _ in range(rng.randrange(0, 6))}


@pytest.mark.parametrize("seed", range(40))
def test_bytes_matches_cpython_json(oracle, seed):
    rng = random.Random(20261106 + seed)
    tree = {"k%d" % i: rand_tree(rng) for i in range(8)}
    # backslash-b / backslash-f must not appear literally in keys for the post-processing above
    raw = json.dumps(tree, ensure_ascii=rng.random() < 0.5).encode()
    s = oracle.New("x", raw)
    assert s.Bytes() == py_go_bytes(tree)
    # idempotence: re-parsing the output and serialising again is a fixed point
    assert oracle.New("x", s.Bytes()).Bytes() == s.Bytes()


def test_root_forms(oracle):
    assert oracle.New("x", b"null").Bytes() == b"{}"         # gabs.go:395-402: nil -> "{}"
    assert oracle.New("x", b" {} ").Bytes() == b"{}"
    assert oracle.New("x", b"[]").Bytes() == b"[]"
    assert oracle.New("x", b'"s"').Bytes() == b'"s"'
    assert oracle.New("x", b"[[],{}]").Bytes() == b"[\n\t[],\n\t{}\n]"
    assert oracle.New("x", b'{"a":{"b":[1]}}').Bytes() == b'{\n\t"a": {\n\t\t"b": [\n\t\t\t1\n\t\t]\n\t}\n}'


STRING_VECTORS = [
    # (JSON text in, Go Marshal text out) -- hand-derived from encoding/json encodeState.string
    (rb'"\u0000\u001f"', rb'"\u0000\u001f"'),
    (rb'"\b\f\n\r\t"', rb'"\u0008\u000c\n\r\t"'),
    (rb'"\/\"\\"', rb'"/\"\\"'),
    (rb'"<>&"', rb'"\u003c\u003e\u0026"'),
    (b'"\x7f"', b'"\x7f"'),
    (rb'"\u00e9"', '"é"'.encode()),
    (rb'"\u2028\u2029"', rb'"\u2028\u2029"'),
    ('"\u2028"'.encode(), rb'"\u2028"'),
    (rb'"\ud83d\ude00"', '"\U0001F600"'.encode()),          # surrogate pair combines
    (rb'"\ud83d"', '"\ufffd"'.encode()),                     # lone high surrogate
    (rb'"\ude00"', '"\ufffd"'.encode()),                     # lone low surrogate
    (rb'"\ud83dx"', '"\ufffdx"'.encode()),
    (rb'"\ud83d\u0041"', '"\ufffdA"'.encode()),              # high + non-surrogate escape
    (rb'"\ud83d\ud83d\ude00"', '"\ufffd\U0001F600"'.encode()),
    (b'"\xff"', '"\ufffd"'.encode()),                        # invalid byte -> U+FFFD at decode time
    (b'"\xe2\x82"', '"\ufffd\ufffd"'.encode()),              # truncated sequence: one per byte
    (b'"\xc0\xaf"', '"\ufffd\ufffd"'.encode()),              # overlong
    (b'"\xed\xa0\x80"', '"\ufffd\ufffd\ufffd"'.encode()),    # UTF-8 encoded surrogate
    (b'"\xf4\x90\x80\x80"', '"\ufffd\ufffd\ufffd\ufffd"'.encode()),  # > U+10FFFF
    (b'"\xf0\x9f\x98\x80"', '"\U0001F600"'.encode()),
]


@pytest.mark.parametrize("src,want", STRING_VECTORS)
def test_string_vectors(oracle, src, want):
    assert oracle.New("x", src).Bytes() == want
    # and as an object key
    assert oracle.New("x", b"{" + src + b":1}").BytesCompact() == b"{" + want + b":1}"


def test_raw_invalid_utf8_in_path_key_is_escaped(oracle):
    # a key that never went through the decoder (it arrives in a path argument) keeps its raw
    # bytes in the tree; Marshal then writes each invalid byte as the six characters \ufffd
    s = oracle.New("x", b"{}")
    s._set(b"a\xffb", b'"v"')
    assert s.BytesCompact() == b'{"a\\ufffdb":"v"}'


def go_float(f: float) -> str:
    """encoding/json floatEncoder restated from CPython's shortest repr (independent of the C code)."""
    if f == 0:
        return "-0" if struct.pack(">d", f)[0] & 0x80 else "0"
    r = repr(abs(f))
    mant, _, exp = r.partition("e")
    ip, _, fp = mant.partition(".")
    e10 = int(exp) if exp else 0
    digits = (ip + fp).lstrip("0")
    # position of the decimal point relative to the first significant digit
    lead = len(ip + fp) - len((ip + fp).lstrip("0"))
    point = len(ip) - lead + e10          # digits[:point] is the integer part
    digits = digits.rstrip("0") or "0"
    sign = "-" if f < 0 else ""
    a = abs(f)
    if a < 1e-6 or a >= 1e21:
        e = point - 1
        m = digits[0] + ("." + digits[1:] if len(digits) > 1 else "")
        return "%s%se%s%s" % (sign, m, "-" if e < 0 else "+", ("%d" % abs(e)) if e < 0 else ("%02d" % e))
    if point <= 0:
        return sign + "0." + "0" * (-point) + digits
    if point >= len(digits):
        return sign + digits + "0" * (point - len(digits))
    return sign + digits[:point] + "." + digits[point:]


FLOAT_VECTORS = [
    ("0", "0"), ("-0", "-0"), ("-0.0", "-0"), ("1", "1"), ("1.0", "1"), ("1.5", "1.5"), ("10", "10"),
    ("100", "100"), ("1e2", "100"), ("1E+2", "100"), ("0.1", "0.1"), ("0.000001", "0.000001"),
    ("0.0000001", "1e-7"), ("1.5e-7", "1.5e-7"), ("1e-10", "1e-10"), ("1e20", "100000000000000000000"),
    ("1e21", "1e+21"), ("1.234e25", "1.234e+25"), ("123456789012345678", "123456789012345680"),
    ("12345678901234567890", "12345678901234567000"), ("0.30000000000000004", "0.30000000000000004"),
    ("3.141592653589793", "3.141592653589793"), ("5e-324", "5e-324"), ("1.7976931348623157e308", "1.7976931348623157e+308"),
    ("-2.5e-8", "-2.5e-8"), ("1e-400", "0"), ("9007199254740993", "9007199254740992"),
]


@pytest.mark.parametrize("src,want", FLOAT_VECTORS)
def test_float_vectors(oracle, src, want):
    assert oracle.New("x", src.encode()).Bytes() == want.encode()
    assert go_float(float(src)) == want


def test_float_random_against_repr_restatement(oracle):
    rng = random.Random(20261106)
    vals = []
    for _ in range(3000):
        bits = rng.getrandbits(64)
        f = struct.unpack(">d", struct.pack(">Q", bits))[0]
        if f != f or f in (float("inf"), float("-inf")):
            continue
        vals.append(f)
    for _ in range(2000):
        vals.append(rng.uniform(-1e6, 1e6))
        vals.append(rng.uniform(-1, 1) * 10 ** rng.randrange(-30, 30))
    raw = ("[" + ",".join(repr(v) for v in vals) + "]").encode()
    got = oracle.New("x", raw).BytesCompact()
    want = "[" + ",".join(go_float(v) for v in vals) + "]"
    assert got.decode() == want


ERROR_VECTORS = [
    # Go error texts restated from encoding/json scanner.go; none is pinned by a reference test
    (b"", "unexpected end of JSON input"),
    (b"   ", "unexpected end of JSON input"),
    (b"{", "unexpected end of JSON input"),
    (b'{"a"', "unexpected end of JSON input"),
    (b'{"a":', "unexpected end of JSON input"),
    (b'"abc', "unexpected end of JSON input"),
    (b"[1", "unexpected end of JSON input"),
    (b'{dfads"', "invalid character 'd' looking for beginning of object key string"),
    (b"{}\n{}", "invalid character '{' after top-level value"),
    (b"[1,]", "invalid character ']' looking for beginning of value"),
    (b'{"a":1,}', "invalid character '}' looking for beginning of object key string"),
    (b'{"a" 1}', "invalid character '1' after object key"),
    (b'{"a":1 "b":2}', "invalid character '\"' after object key:value pair"),
    (b"[1 2]", "invalid character '2' after array element"),
    (b"'a'", "invalid character '\\'' looking for beginning of value"),
    (b'"\x01"', "invalid character '\\x01' in string literal"),
    (b'"\n"', "invalid character '\\n' in string literal"),
    (b'"\\x"', "invalid character 'x' in string escape code"),
    (b'"\\u12g4"', "invalid character 'g' in \\u hexadecimal character escape"),
    (b"-x", "invalid character 'x' in numeric literal"),
    (b"01", "invalid character '1' after top-level value"),
    (b"1.x", "invalid character 'x' after decimal point in numeric literal"),
    (b"1e+x", "invalid character 'x' in exponent of numeric literal"),
    (b"tru", "invalid character ' ' in literal true (expecting 'e')"),
    (b"trux", "invalid character 'x' in literal true (expecting 'e')"),
    (b"fals", "invalid character ' ' in literal false (expecting 'e')"),
    (b"nul", "invalid character ' ' in literal null (expecting 'l')"),
    (b"\xe9", "invalid character '\u00e9' looking for beginning of value"),
    (b"\x80", "invalid character '\\u0080' looking for beginning of value"),
    (b'{"a":1e999}', "json: cannot unmarshal number 1e999 into Go value of type float64"),
    (b"[" * 10001 + b"]" * 10001, "exceeded max depth"),
]


@pytest.mark.parametrize("src,msg", ERROR_VECTORS)
def test_error_vectors(oracle, src, msg):
    with pytest.raises(OracleError) as e:
        oracle.New("x", src)
    assert str(e.value) == msg


def test_depth_limit_boundary(oracle):
    assert oracle.New("x", b"[" * 10000 + b"]" * 10000).BytesCompact() == b"[" * 10000 + b"]" * 10000


def test_duplicate_keys_last_wins(oracle):
    assert oracle.New("x", b'{"a":1,"b":2,"a":3}').BytesCompact() == b'{"a":3,"b":2}'
    assert oracle.New("x", b'{"a":{"x":1},"a":"s"}').BytesCompact() == b'{"a":"s"}'
    # keys that only become equal after decoding
    assert oracle.New("x", b'{"\\u0061":1,"a":2}').BytesCompact() == b'{"a":2}'
    assert oracle.New("x", b'{"\xff":1,"\xfe":2}').BytesCompact() == '{"\ufffd":2}'.encode()


def test_key_order_is_bytewise(oracle):
    doc = '{"b":1,"a":1,"B":1,"é":1,"aa":1,"":1,"a\\u0000":1,"~":1,"\U0001F600":1,"\uffff":1}'.encode()
    keys = ["", "B", "a", "a\x00", "aa", "b", "~", "é", "\uffff", "\U0001F600"]
    assert keys == sorted(keys, key=lambda k: k.encode())
    got = json.loads(oracle.New("x", doc).BytesCompact(), object_pairs_hook=list)
    assert [k for k, _ in got] == keys
