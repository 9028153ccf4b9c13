// COLMAP text export: the text the reference builds with Python string `+=` in per-row loops
// (dpvo/plot_utils.py:82-93: images.txt, points3D.txt), written in one pass into a caller-provided buffer.  Host code only.
// Every number goes through the same formatting the reference's f-strings / str() apply: CPython's float repr (shortest digits
// that round-trip; fixed notation while -4 < decimal exponent <= 16, else d.ddde+XX with at least two exponent digits).
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstring>

#include "../../include/wvb200.h"

namespace {

// repr(float) of CPython (Python/pystrtod.c format_float_short, mode 'r', Py_DTSF_ADD_DOT_0); returns the length
int py_float_repr(double v, char* out) {
  if (std::isnan(v)) { memcpy(out, "nan", 3); return 3; }
  if (std::isinf(v)) { if (v < 0) { memcpy(out, "-inf", 4); return 4; } memcpy(out, "inf", 3); return 3; }
  char sci[40];
  const auto r = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);  // [-]d[.ddd]e[+-]XX, shortest round-trip
  char digits[24];
  int nd = 0, n = 0;
  const char* p = sci;
  if (*p == '-') { out[n++] = '-'; p++; }
  for (; p < r.ptr && *p != 'e'; p++) if (*p != '.') digits[nd++] = *p;
  int ex = 0;
  { const char* q = p + 1; const bool neg = *q == '-'; q++; for (; q < r.ptr; q++) ex = ex * 10 + (*q - '0'); if (neg) ex = -ex; }
  const int decpt = ex + 1;
  if (decpt > -4 && decpt <= 16) {
    if (decpt <= 0) {
      out[n++] = '0'; out[n++] = '.';
      for (int i = 0; i < -decpt; i++) out[n++] = '0';
      memcpy(out + n, digits, nd); n += nd;
    } else if (decpt >= nd) {
      memcpy(out + n, digits, nd); n += nd;
      for (int i = nd; i < decpt; i++) out[n++] = '0';
      out[n++] = '.'; out[n++] = '0';
    } else {
      memcpy(out + n, digits, decpt); n += decpt;
      out[n++] = '.';
      memcpy(out + n, digits + decpt, nd - decpt); n += nd - decpt;
    }
    return n;
  }
  out[n++] = digits[0];
  if (nd > 1) { out[n++] = '.'; memcpy(out + n, digits + 1, nd - 1); n += nd - 1; }
  out[n++] = 'e';
  int e = decpt - 1;
  out[n++] = e < 0 ? '-' : '+';
  if (e < 0) e = -e;
  char eb[8];
  int ne = 0;
  do { eb[ne++] = (char)('0' + e % 10); e /= 10; } while (e);
  if (ne < 2) eb[ne++] = '0';
  while (ne) out[n++] = eb[--ne];
  return n;
}

int put_int(long long v, char* out) {
  const auto r = std::to_chars(out, out + 24, v);
  return (int)(r.ptr - out);
}

}  // namespace

extern "C" {

int wv_format_f64(double v, char* out32) { return py_float_repr(v, out32); }

// points3D.txt (dpvo/plot_utils.py:88-93): "{i} {x} {y} {z} {r} {g} {b} 0.0 0 0 0 0 0 0\n", i = first_id ..
int64_t wv_colmap_points3d_text(const double* xyz, const uint8_t* rgb, int64_t n, int64_t first_id, char* out, int64_t cap) {
  if (n < 0 || (n > 0 && (!xyz || !rgb)) || !out) return WV_EINVAL;
  static const char tail[] = " 0.0 0 0 0 0 0 0\n";
  int64_t w = 0;
  for (int64_t i = 0; i < n; i++) {
    if (cap - w < 160) return WV_EWORKSPACE;  // a row is at most 20 + 3 * 25 + 3 * 4 + 17 bytes
    w += put_int(first_id + i, out + w);
    for (int k = 0; k < 3; k++) { out[w++] = ' '; w += py_float_repr(xyz[3 * i + k], out + w); }
    for (int k = 0; k < 3; k++) { out[w++] = ' '; w += put_int(rgb[3 * i + k], out + w); }
    memcpy(out + w, tail, sizeof(tail) - 1); w += sizeof(tail) - 1;
  }
  return w;
}

// images.txt (dpvo/plot_utils.py:82-87): "{idx} {qw} {qx} {qy} {qz} {x} {y} {z} 1 {name}\n\n", idx = 1 ..;
// pose rows are (qw, qx, qy, qz, x, y, z); names: n NUL-terminated strings back to back
int64_t wv_colmap_images_text(const double* pose7, const char* names, int64_t n, char* out, int64_t cap) {
  if (n < 0 || (n > 0 && (!pose7 || !names)) || !out) return WV_EINVAL;
  int64_t w = 0;
  const char* nm = names;
  for (int64_t i = 0; i < n; i++) {
    const size_t ln = strlen(nm);
    if (cap - w < (int64_t)(240 + ln)) return WV_EWORKSPACE;
    w += put_int(i + 1, out + w);
    for (int k = 0; k < 7; k++) { out[w++] = ' '; w += py_float_repr(pose7[7 * i + k], out + w); }
    out[w++] = ' '; out[w++] = '1'; out[w++] = ' ';
    memcpy(out + w, nm, ln); w += ln;
    out[w++] = '\n'; out[w++] = '\n';
    nm += ln + 1;
  }
  return w;
}

}  // extern "C"
