// Result sink of the explanation path: attribution rows as the JSON text the reference stores.
//
// Reference (api/worker.py:82-88): after EXPL.shap_values(X) the task builds
//     'shap_values': json.dumps(shap_values[0].tolist())
// and inserts that string into Postgres.  json.dumps writes a float as float.__repr__ does: the shortest decimal string
// that reads back to the same double, fixed notation when the decimal point position is in (-4, 16], exponent notation
// ("1e-05", "1.5e+16") otherwise, "NaN" / "Infinity" / "-Infinity" for non-finite values, ", " between elements.
// std::to_chars (C++17) yields the same shortest digits; this file lays them out the Python way.  Host code only: a batch
// of attribution rows comes back from the device once and is formatted here at ~1 us per row and thread instead of
// ~10 us per row in the interpreter; rows are independent, so callers may split a batch over threads.

#include <charconv>
#include <cmath>
#include <cstring>

#include "fdb_internal.cuh"

namespace {

constexpr int kMaxFloatChars = 26;        // "-2.2250738585072014e-308" is 24 characters

// repr(float) into out (at least kMaxFloatChars bytes); returns the end pointer
char* py_float_repr(double v, char* out) {
    if (std::isnan(v)) { std::memcpy(out, "NaN", 3); return out + 3; }
    if (std::isinf(v)) {
        if (v < 0) { std::memcpy(out, "-Infinity", 9); return out + 9; }
        std::memcpy(out, "Infinity", 8);
        return out + 8;
    }
    char sci[40];
    auto res = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);      // shortest round trip: [-]d[.ddd]e[+-]XX
    const char* p = sci;
    if (*p == '-') *out++ = *p++;
    const char* e = p;
    while (*e != 'e') ++e;
    char digits[24];
    int nd = 0;
    for (const char* q = p; q < e; ++q)
        if (*q != '.') digits[nd++] = *q;
    int exp10 = 0;
    {
        const char* q = e + 1;
        const bool neg = *q == '-';
        ++q;                                        // sign is always written in scientific format
        for (; q < res.ptr; ++q) exp10 = exp10 * 10 + (*q - '0');
        if (neg) exp10 = -exp10;
    }
    if (exp10 < -4 || exp10 >= 16) {                // exponent notation, exactly what to_chars wrote
        const size_t len = (size_t)(res.ptr - p);
        std::memcpy(out, p, len);
        return out + len;
    }
    if (exp10 >= 0) {                               // d...d.d...d, ".0" when nothing is left after the point
        for (int i = 0; i <= exp10; ++i) *out++ = i < nd ? digits[i] : '0';
        *out++ = '.';
        if (nd > exp10 + 1)
            for (int i = exp10 + 1; i < nd; ++i) *out++ = digits[i];
        else
            *out++ = '0';
        return out;
    }
    *out++ = '0';                                   // 0.000ddd
    *out++ = '.';
    for (int i = 0; i < -exp10 - 1; ++i) *out++ = '0';
    for (int i = 0; i < nd; ++i) *out++ = digits[i];
    return out;
}

}  // namespace

// json.dumps(row.tolist()) for every row of a (n_rows, n_cols) float64 array: the texts are written back to back into
// `out` (no terminators), row r occupies [offsets[r], offsets[r + 1]).  out_cap >= n_rows * (n_cols * 28 + 2) always fits.
extern "C" int fdb_format_json_rows(const double* values, int64_t n_rows, int n_cols, char* out, int64_t out_cap, int64_t* offsets) {
    FDB_REQUIRE(n_rows >= 0 && n_cols >= 0, FDB_ERR_ARG, "negative size");
    FDB_REQUIRE(offsets != nullptr && (n_rows == 0 || (values != nullptr && out != nullptr)), FDB_ERR_ARG, "NULL argument");
    const int64_t row_cap = (int64_t)n_cols * (kMaxFloatChars + 2) + 2;
    char* w = out;
    offsets[0] = 0;
    for (int64_t r = 0; r < n_rows; ++r) {
        FDB_REQUIRE((w - out) + row_cap <= out_cap, FDB_ERR_ARG, "output buffer too small: %lld bytes for %lld rows of %d values",
                    (long long)out_cap, (long long)n_rows, n_cols);
        const double* row = values + r * n_cols;
        *w++ = '[';
        for (int c = 0; c < n_cols; ++c) {
            if (c) { *w++ = ','; *w++ = ' '; }
            w = py_float_repr(row[c], w);
        }
        *w++ = ']';
        offsets[r + 1] = w - out;
    }
    return FDB_OK;
}
