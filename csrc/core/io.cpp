// File formats of libhpnn: nn.conf, kernel files, sample files, directory walks.
// Behavioural reference (not code): src/libhpnn.c:658-937 (conf), src/ann.c:206-631
// (kernel load), src/ann.c:770-857 (kernel dump), src/libhpnn.c:1070-1145 (samples),
// src/libhpnn.c:1184-1229 (directory listing + shuffle).
#include <dirent.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <charconv>
#include <time.h>

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <string>
#include <vector>

#include "hpnn_core.h"
#include "kernel_host.hpp"

namespace {

char *dup_cstr(const std::string &s) {
  char *p = (char *)malloc(s.size() + 1);
  if (p) memcpy(p, s.c_str(), s.size() + 1);
  return p;
}

// value part of "[keyword...] value": everything after the closing bracket,
// leading blanks dropped, cut at the first blank / '#' / newline.
std::string bracket_value(const char *after_key) {
  const char *p = strchr(after_key, ']');
  p = p ? p + 1 : after_key;
  while (*p && !isgraph((unsigned char)*p)) ++p;
  const char *e = p;
  while (*e && isgraph((unsigned char)*e) && *e != '#') ++e;
  return std::string(p, e);
}

const char *after_bracket(const char *key_pos) {
  const char *p = strchr(key_pos, ']');
  p = p ? p + 1 : key_pos;
  while (*p && !isgraph((unsigned char)*p)) ++p;
  return p;
}

bool read_line(FILE *fp, std::string &out) {
  static thread_local char *buf = nullptr;
  static thread_local size_t cap = 0;
  ssize_t n = getline(&buf, &cap, fp);
  if (n < 0) return false;
  out.assign(buf, (size_t)n);
  return true;
}

// parse up to `want` doubles from a text line; returns how many were read
int parse_doubles(const char *s, double *dst, int want) {
  int got = 0;
  char *end = nullptr;
  const char *stop = s + strlen(s);
  while (got < want) {
    while (*s && !isgraph((unsigned char)*s)) ++s;
    if (!*s) break;
    // std::from_chars is exactly rounded like strtod and several times faster; whatever it does not
    // take (leading '+', hex floats, "inf", ...) goes through strtod as before
    double v;
    const auto r = std::from_chars(s, stop, v);
    if (r.ec == std::errc() && r.ptr != s) {
      end = const_cast<char *>(r.ptr);
    } else {
      v = strtod(s, &end);
      if (end == s) break;
    }
    dst[got++] = v;
    s = end;
  }
  return got;
}

std::vector<long> parse_uints(const char *s) {
  std::vector<long> v;
  while (*s) {
    while (*s && !isdigit((unsigned char)*s)) {
      if (*s == '#') return v;
      ++s;
    }
    if (!*s) break;
    char *end;
    v.push_back((long)strtoull(s, &end, 10));
    s = end;
  }
  return v;
}

}  // namespace

extern "C" {

void hpnn_free(void *p) { free(p); }

/* ------------------------------ kernel object ---------------------------- */
hpnn_kernel *hpnn_kernel_create(int n_dims, const int *dims) {
  if (n_dims < 3 || n_dims > (1 << 16) || !dims) return nullptr;  // any depth: only device nets are capped
  for (int i = 0; i < n_dims; ++i)
    if (dims[i] < 1) return nullptr;
  hpnn_kernel *k = new hpnn_kernel();
  k->dims.assign(dims, dims + n_dims);
  k->W.resize(n_dims - 1);
  for (int l = 0; l < n_dims - 1; ++l) k->W[l].assign((size_t)dims[l + 1] * dims[l], 0.0);
  return k;
}
void hpnn_kernel_destroy(hpnn_kernel *k) { delete k; }
int hpnn_kernel_n_layers(const hpnn_kernel *k) { return k ? k->n_layers() : 0; }
int hpnn_kernel_dim(const hpnn_kernel *k, int i) {
  return (k && i >= 0 && i < (int)k->dims.size()) ? k->dims[i] : 0;
}
double *hpnn_kernel_weights(hpnn_kernel *k, int l) {
  return (k && l >= 0 && l < k->n_layers()) ? k->W[l].data() : nullptr;
}
const char *hpnn_kernel_name(const hpnn_kernel *k) { return (k && k->has_name) ? k->name.c_str() : nullptr; }
void hpnn_kernel_set_name(hpnn_kernel *k, const char *name) {
  if (!k) return;
  k->has_name = name != nullptr;
  k->name = name ? name : "";
}

hpnn_kernel *hpnn_kernel_generate(unsigned *seed, int n_dims, const int *dims) {
  hpnn_kernel *k = hpnn_kernel_create(n_dims, dims);
  if (!k) return nullptr;
  unsigned s = seed ? *seed : 0;
  if (s == 0) s = (unsigned)time(nullptr);
  if (seed) *seed = s;
  srandom(s);
  // one stream of glibc random() draws, layer by layer, row-major
  for (int l = 0; l < k->n_layers(); ++l) {
    const double inv = sqrt((double)k->dims[l]);
    for (double &w : k->W[l]) {
      double r = (double)random() / RAND_MAX;
      w = 2.0 * (r - 0.5) / inv;
    }
  }
  return k;  // generated kernels stay unnamed (dumped as "(null)")
}

hpnn_kernel *hpnn_kernel_load(const char *path) {
  FILE *fp = path ? fopen(path, "r") : nullptr;
  if (!fp) {
    fprintf(stderr, "NN(ERR): Error opening kernel file: %s\n", path ? path : "(null)");
    return nullptr;
  }
  std::string line;
  hpnn_kernel *k = nullptr;
  std::string name;
  auto fail = [&](const char *msg) -> hpnn_kernel * {
    fprintf(stderr, "NN(ERR): kernel read: %s\n", msg);
    fclose(fp);
    delete k;
    return nullptr;
  };
  if (!read_line(fp, line) || line.find("[name]") == std::string::npos)
    return fail("kernel file should start with [name] keyword!");
  {
    const char *p = after_bracket(line.c_str() + line.find("[name]"));
    name = p;
    while (!name.empty() && (name.back() == '\n' || name.back() == '\r')) name.pop_back();
    if (name.empty()) name = "noname";
  }
  // topology
  std::vector<int> dims;
  while (read_line(fp, line)) {
    size_t pos = line.find("[param]");
    if (pos == std::string::npos) continue;
    for (long v : parse_uints(line.c_str() + pos + 7)) dims.push_back((int)v);
    break;
  }
  if (dims.size() < 3) return fail("parameter line has too few parameters!");
  for (int d : dims)
    if (d < 1) return fail("zero in parameter line!");
  k = hpnn_kernel_create((int)dims.size(), dims.data());
  if (!k) return fail("allocation failed");
  hpnn_kernel_set_name(k, name.c_str());
  const int L = k->n_layers();  // last one is the output layer
  std::vector<char> seen(L, 0);
  // layer blocks, in any order
  bool have_line = read_line(fp, line);
  while (have_line) {
    int layer = -1;
    size_t pos;
    if ((pos = line.find("[hidden")) != std::string::npos) {
      std::vector<long> v = parse_uints(line.c_str() + pos);
      if (v.size() < 2 || v[0] < 1 || v[0] > L - 1) return fail("malformed hidden layer definition");
      layer = (int)v[0] - 1;
      if (v[1] != dims[layer + 1]) return fail("inconsistent neuron number!");
    } else if ((pos = line.find("[output]")) != std::string::npos) {
      std::vector<long> v = parse_uints(line.c_str() + pos);
      if (v.empty() || v[0] != dims[L]) return fail("inconsistent neuron number for output!");
      layer = L - 1;
    }
    if (layer < 0) {
      have_line = read_line(fp, line);
      continue;
    }
    const int N = dims[layer + 1], M = dims[layer];
    double *W = k->W[layer].data();
    for (int j = 0; j < N; ++j) {
      if (!read_line(fp, line) || line.find("[neuron") == std::string::npos)
        return fail("neuron definition missing!");
      std::vector<long> v = parse_uints(line.c_str() + line.find("[neuron"));
      if (v.size() < 2 || v[0] < 1 || v[1] < 1) return fail("malformed neuron line");
      if (v[1] > M) return fail("neuron inconsistent input number!");
      if (!read_line(fp, line)) return fail("missing weight line");
      if (parse_doubles(line.c_str(), W + (size_t)j * M, (int)v[1]) != (int)v[1])
        return fail("short weight line");
    }
    seen[layer] = 1;
    have_line = read_line(fp, line);
  }
  fclose(fp);
  return k;  // layers without a block keep zero weights, as in the reference
}

// "%17.15f" through std::to_chars: the same correctly rounded digits as printf (checked on 2 M
// values incl. signed zeros and 1e300), three times faster; the kernel file of the 784-300-10
// tutorial net has 238 000 of them and train_nn writes it twice.
static inline char *fmt_weight(char *p, double x) {
  char tmp[352];  // DBL_MAX has 309 integer digits
  const auto r = std::to_chars(tmp, tmp + sizeof(tmp), x, std::chars_format::fixed, 15);
  const int n = (int)(r.ptr - tmp);
  for (int i = n; i < 17; ++i) *p++ = ' ';
  memcpy(p, tmp, (size_t)n);
  return p + n;
}

static void dump_layer(FILE *fp, const double *W, int N, int M, std::vector<char> &buf) {
  const size_t cap = (size_t)1 << 16;
  if (buf.size() < cap + 512) buf.resize(cap + 512);
  for (int j = 0; j < N; ++j) {
    fprintf(fp, "[neuron %i] %i\n", j + 1, M);
    char *p = buf.data();
    const double *w = W + (size_t)j * M;
    for (int c = 0; c < M; ++c) {
      if (c) *p++ = ' ';
      p = fmt_weight(p, w[c]);
      if ((size_t)(p - buf.data()) > cap) {  // any magnitude fits: flush in pieces
        fwrite(buf.data(), 1, (size_t)(p - buf.data()), fp);
        p = buf.data();
      }
    }
    *p++ = '\n';
    fwrite(buf.data(), 1, (size_t)(p - buf.data()), fp);
  }
}

int hpnn_kernel_dump(const hpnn_kernel *k, FILE *fp) {
  if (!k || !fp) return -1;
  const int L = k->n_layers();
  if (k->has_name) fprintf(fp, "[name] %s\n", k->name.c_str());
  else fprintf(fp, "[name] (null)\n");
  fprintf(fp, "[param] %i", k->dims[0]);
  for (int l = 1; l <= L; ++l) fprintf(fp, " %i", k->dims[l]);
  fprintf(fp, "\n[input] %i\n", k->dims[0]);
  std::vector<char> buf;
  for (int l = 0; l < L - 1; ++l) {
    fprintf(fp, "[hidden %i] %i\n", l + 1, k->dims[l + 1]);
    dump_layer(fp, k->W[l].data(), k->dims[l + 1], k->dims[l], buf);
  }
  fprintf(fp, "[output] %i\n", k->dims[L]);
  dump_layer(fp, k->W[L - 1].data(), k->dims[L], k->dims[L - 1], buf);
  return ferror(fp) ? -1 : 0;
}

int hpnn_kernel_dump_path(const hpnn_kernel *k, const char *path) {
  FILE *fp = fopen(path, "w");
  if (!fp) return -1;
  int rc = hpnn_kernel_dump(k, fp);
  fclose(fp);
  return rc;
}

static const char BIN_MAGIC[8] = {'H', 'P', 'N', 'N', 'B', '2', '0', '0'};

int hpnn_kernel_dump_binary(const hpnn_kernel *k, const char *path) {
  FILE *fp = (k && path) ? fopen(path, "wb") : nullptr;
  if (!fp) return -1;
  int32_t nd = (int32_t)k->dims.size(), nl = k->has_name ? (int32_t)k->name.size() : -1;
  fwrite(BIN_MAGIC, 1, 8, fp);
  fwrite(&nd, 4, 1, fp);
  fwrite(k->dims.data(), 4, nd, fp);
  fwrite(&nl, 4, 1, fp);
  if (nl > 0) fwrite(k->name.data(), 1, nl, fp);
  for (auto &w : k->W) fwrite(w.data(), 8, w.size(), fp);
  int rc = ferror(fp) ? -1 : 0;
  fclose(fp);
  return rc;
}

hpnn_kernel *hpnn_kernel_load_binary(const char *path) {
  FILE *fp = path ? fopen(path, "rb") : nullptr;
  if (!fp) return nullptr;
  char magic[8];
  int32_t nd = 0, nl = 0;
  hpnn_kernel *k = nullptr;
  std::vector<int> dims;
  bool ok = fread(magic, 1, 8, fp) == 8 && !memcmp(magic, BIN_MAGIC, 8) && fread(&nd, 4, 1, fp) == 1 &&
            nd >= 3 && nd <= (1 << 16);
  if (ok) {
    dims.resize(nd);
    ok = fread(dims.data(), 4, nd, fp) == (size_t)nd && fread(&nl, 4, 1, fp) == 1 && nl < (1 << 20);
  }
  if (ok) {
    std::string name;
    if (nl > 0) {
      name.resize(nl);
      ok = fread(&name[0], 1, nl, fp) == (size_t)nl;
    }
    if (ok) k = hpnn_kernel_create(nd, dims.data());
    if (k && nl >= 0) hpnn_kernel_set_name(k, name.c_str());
  }
  if (k)
    for (auto &w : k->W)
      if (fread(w.data(), 8, w.size(), fp) != w.size()) {
        delete k;
        k = nullptr;
        break;
      }
  fclose(fp);
  return k;
}

/* --------------------------------- nn.conf ------------------------------- */
void hpnn_conf_free(hpnn_conf *c) {
  if (!c) return;
  free(c->name);
  free(c->f_kernel);
  free(c->samples);
  free(c->tests);
  free(c->hid);
  memset(c, 0, sizeof(*c));
  c->type = HPNN_TYPE_UKN;
  c->train = HPNN_TRAIN_UKN;
}

int hpnn_conf_load(const char *path, hpnn_conf *c) {
  if (!c) return -1;
  memset(c, 0, sizeof(*c));
  c->type = HPNN_TYPE_UKN;
  c->train = HPNN_TRAIN_UKN;
  FILE *fp = path ? fopen(path, "r") : nullptr;
  if (!fp) {
    fprintf(stderr, "NN(ERR): Error opening configuration file: %s\n", path ? path : "(null)");
    return -1;
  }
  std::string line;
  bool bad = false;
  auto malformed = [&](const char *key, const char *val) {
    fprintf(stderr, "NN(ERR): Malformed NN configuration file!\n");
    fprintf(stderr, "NN(ERR): [%s] value: %s\n", key, val);
    bad = true;
  };
  while (!bad && read_line(fp, line)) {
    const char *s = line.c_str();
    const char *p;
    if ((p = strstr(s, "[name"))) {
      free(c->name);
      std::string v = bracket_value(p);
      c->name = v.empty() ? nullptr : dup_cstr(v);
    }
    if ((p = strstr(s, "[type"))) {
      switch (*after_bracket(p)) {
        case 'L': c->type = HPNN_TYPE_LNN; break;
        case 'S': c->type = HPNN_TYPE_SNN; break;
        default: c->type = HPNN_TYPE_ANN; break;
      }
    }
    if ((p = strstr(s, "[init"))) {
      if (strstr(s, "generate") || strstr(s, "GENERATE")) {
        c->need_init = 1;
      } else {
        c->need_init = 0;
        std::string v = bracket_value(p);
        if (v.empty()) {
          fprintf(stderr, "NN(ERR): Malformed NN configuration file!\n");
          fprintf(stderr, "NN(ERR): [init] can't read filename: %s\n", after_bracket(p));
          bad = true;
        } else {
          free(c->f_kernel);
          c->f_kernel = dup_cstr(v);
        }
      }
    }
    if ((p = strstr(s, "[seed"))) {
      const char *v = after_bracket(p);
      if (!isdigit((unsigned char)*v)) malformed("seed", v);
      else c->seed = (unsigned)strtoull(v, nullptr, 10);
    }
    if ((p = strstr(s, "[input"))) {
      const char *v = after_bracket(p);
      if (!isdigit((unsigned char)*v)) malformed("input", v);
      else c->n_in = (int)strtoull(v, nullptr, 10);
    }
    if ((p = strstr(s, "[hidden"))) {
      const char *v = after_bracket(p);
      if (!isdigit((unsigned char)*v)) malformed("hidden", v);
      else {
        std::vector<long> h = parse_uints(v);
        free(c->hid);
        c->n_hid = (int)h.size();
        c->hid = (int *)malloc(sizeof(int) * h.size());
        for (size_t i = 0; i < h.size(); ++i) c->hid[i] = (int)h[i];
      }
    }
    if ((p = strstr(s, "[output"))) {
      const char *v = after_bracket(p);
      if (!isdigit((unsigned char)*v)) malformed("output", v);
      else c->n_out = (int)strtoull(v, nullptr, 10);
    }
    if ((p = strstr(s, "[train"))) {
      const char *v = after_bracket(p);
      switch (*v) {
        case 'B': c->train = (v[1] && v[2] == 'M') ? HPNN_TRAIN_BPM : HPNN_TRAIN_BP; break;
        case 'C': c->train = HPNN_TRAIN_CG; break;
        case 'S': c->train = HPNN_TRAIN_SPLX; break;
        default: c->train = HPNN_TRAIN_UKN;
      }
    }
    if ((p = strstr(s, "[sample_dir"))) {
      free(c->samples);
      std::string v = bracket_value(p);
      c->samples = v.empty() ? nullptr : dup_cstr(v);
    }
    if ((p = strstr(s, "[test_dir"))) {
      free(c->tests);
      std::string v = bracket_value(p);
      c->tests = v.empty() ? nullptr : dup_cstr(v);
    }
  }
  fclose(fp);
  if (!bad && c->type == HPNN_TYPE_UKN) {
    fprintf(stderr, "NN(ERR): Malformed NN configuration file!\nNN(ERR): [type] unknown or missing...\n");
    bad = true;
  }
  if (!bad && c->need_init) {
    const char *miss = nullptr;
    if (c->n_in == 0) miss = "input";
    else if (c->n_hid == 0) miss = "hidden";
    else if (c->n_out == 0) miss = "output";
    if (miss) {
      fprintf(stderr, "NN(ERR): Malformed NN configuration file!\nNN(ERR): [%s] wrong or missing...\n", miss);
      bad = true;
    }
    for (int i = 0; !bad && i < c->n_hid; ++i)
      if (c->hid[i] == 0) {
        fprintf(stderr, "NN(ERR): Malformed NN configuration file!\n");
        fprintf(stderr, "NN(ERR): [hidden] some have a 0 neuron content!\n");
        bad = true;
      }
  }
  if (bad) {
    hpnn_conf_free(c);
    return -1;
  }
  return 0;
}

int hpnn_conf_dump(const hpnn_conf *c, int n_dims, const int *dims, FILE *fp) {
  if (!c || !fp) return -1;
  fprintf(fp, "# NN configuration\n");
  if (c->name) fprintf(fp, "[name] %s\n", c->name);
  else fprintf(fp, "[name] (null)\n");
  fprintf(fp, "[type] %s\n", c->type == HPNN_TYPE_LNN ? "LNN" : c->type == HPNN_TYPE_SNN ? "SNN" : "ANN");
  if (c->need_init) fprintf(fp, "[init] generate\n");
  else if (c->f_kernel) fprintf(fp, "[init] %s\n", c->f_kernel);
  else fprintf(fp, "[init] INVALID <- this should trigger an error\n");
  fprintf(fp, "[seed] %i\n", (int)c->seed);
  fprintf(fp, "[inputs] %i\n", (dims && n_dims > 0) ? dims[0] : 0);
  fprintf(fp, "[hiddens] ");
  for (int i = 1; dims && i < n_dims - 1; ++i) fprintf(fp, "%i ", dims[i]);
  fprintf(fp, "\n[outputs] %i\n", (dims && n_dims > 1) ? dims[n_dims - 1] : 0);
  switch (c->train) {
    case HPNN_TRAIN_BP: fprintf(fp, "[train] BP\n"); break;
    case HPNN_TRAIN_BPM: fprintf(fp, "[train] BPM\n"); break;
    case HPNN_TRAIN_CG: fprintf(fp, "[train] CG\n"); break;
    case HPNN_TRAIN_SPLX: fprintf(fp, "[train] SPLX\n"); break;
    default: fprintf(fp, "[train] none\n");
  }
  if (c->samples) fprintf(fp, "[sample_dir] %s\n", c->samples);
  else fprintf(fp, "[sample_dir] INVALID <- this should trigger an error\n");
  if (c->tests) fprintf(fp, "[test_dir] %s\n", c->tests);
  else fprintf(fp, "[test_dir] INVALID <- this should trigger an error\n");
  return 0;
}

/* --------------------------------- samples ------------------------------- */
int hpnn_sample_read(const char *path, double **in, int *n_in, double **out, int *n_out) {
  if (in) *in = nullptr;
  if (out) *out = nullptr;
  if (n_in) *n_in = 0;
  if (n_out) *n_out = 0;
  FILE *fp = path ? fopen(path, "r") : nullptr;
  if (!fp || !in || !out) {
    if (fp) fclose(fp);
    return -1;
  }
  std::string line;
  int rc = 0;
  while (rc == 0 && read_line(fp, line)) {
    for (int which = 0; which < 2 && rc == 0; ++which) {
      const char *key = which ? "[output" : "[input";
      const char *p = strstr(line.c_str(), key);
      if (!p) continue;
      const char *v = after_bracket(p);
      long n = isdigit((unsigned char)*v) ? (long)strtoull(v, nullptr, 10) : 0;
      if (n <= 0) {
        fprintf(stderr, "NN(ERR): sample %s %s read failed!\n", path, which ? "output" : "input");
        rc = -1;
        break;
      }
      double *buf = (double *)calloc((size_t)n, sizeof(double));
      if (!read_line(fp, line) || parse_doubles(line.c_str(), buf, (int)n) < (int)n - 1) {
        // the reference tolerates a missing last value (reads it as 0) but nothing shorter
        free(buf);
        rc = -1;
        break;
      }
      double **dst = which ? out : in;
      free(*dst);
      *dst = buf;
      *(which ? n_out : n_in) = (int)n;
    }
  }
  fclose(fp);
  if (rc != 0) {
    free(*in);
    free(*out);
    *in = *out = nullptr;
  }
  return rc;
}

// Parse many sample files in parallel (OpenMP) straight into two dense row-major matrices:
// the data-loader behind SampleSet / the pipelined training pass.  ok[i] = 1 when file i was
// read and has exactly n_in / n_out values; returns the number of good files.
int hpnn_sample_read_many(const char *const *paths, int n, int n_in, int n_out, double *X, double *T, unsigned char *ok) {
  int good = 0;
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : good)
  for (int i = 0; i < n; ++i) {
    double *in = nullptr, *out = nullptr;
    int ni = 0, no = 0;
    const int rc = hpnn_sample_read(paths[i], &in, &ni, &out, &no);
    const bool fine = rc == 0 && ni == n_in && no == n_out;
    if (fine) {
      memcpy(X + (size_t)i * n_in, in, sizeof(double) * n_in);
      memcpy(T + (size_t)i * n_out, out, sizeof(double) * n_out);
      ++good;
    }
    if (ok) ok[i] = fine ? 1 : 0;
    free(in);
    free(out);
  }
  return good;
}

int hpnn_sample_write(const char *path, const double *in, int n_in, const double *out, int n_out,
                      const char *comment) {
  FILE *fp = path ? fopen(path, "w") : nullptr;
  if (!fp) return -1;
  fprintf(fp, "[input] %i\n", n_in);
  for (int i = 0; i < n_in; ++i) fprintf(fp, i ? " %7.5f" : "%7.5f", in[i]);
  if (comment) fprintf(fp, "\n[output] %i  #%s\n", n_out, comment);
  else fprintf(fp, "\n[output] %i\n", n_out);
  for (int i = 0; i < n_out; ++i) fprintf(fp, i ? " %.1f" : "%.1f", out[i]);
  fprintf(fp, "\n");
  fclose(fp);
  return 0;
}

int hpnn_dir_list(const char *dir, char ***names, int *n) {
  if (names) *names = nullptr;
  if (n) *n = 0;
  DIR *d = dir ? opendir(dir) : nullptr;
  if (!d || !names || !n) {
    if (d) closedir(d);
    return -1;
  }
  std::vector<char *> v;
  while (struct dirent *e = readdir(d)) {
    if (e->d_name[0] == '.') continue;  // also drops "." and ".."
    v.push_back(strdup(e->d_name));
  }
  closedir(d);
  *n = (int)v.size();
  *names = (char **)malloc(sizeof(char *) * (v.size() + 1));
  for (size_t i = 0; i < v.size(); ++i) (*names)[i] = v[i];
  (*names)[v.size()] = nullptr;
  return 0;
}

void hpnn_dir_list_free(char **names, int n) {
  if (!names) return;
  for (int i = 0; i < n; ++i) free(names[i]);
  free(names);
}

void hpnn_shuffle_order(unsigned seed, int n, int *order) {
  if (n <= 0 || !order) return;
  std::vector<char> used(n, 0);
  srandom(seed);
  for (int j = 0; j < n; ++j) {
    unsigned idx;
    do {
      idx = (unsigned)((double)random() * n / RAND_MAX);
    } while (idx >= (unsigned)n || used[idx]);  // idx==n only when random()==RAND_MAX
    used[idx] = 1;
    order[j] = (int)idx;
  }
}

}  // extern "C"
