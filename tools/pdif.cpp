// pdif: RRUFF powder-XRD database (dif/ + raw/ directories) -> hpnn text samples.
//
//   pdif rruff_dir -i N_BINS -o N_OUT [-s samples_dir]
//
// For every DIF file the matching raw XY pattern is integrated between 2theta = 5 and 90
// degrees into N_BINS bins, normalised by the largest bin and prefixed with T/273.15; the
// target is a +-1 one-hot of the space-group number (1..230).  Output format and selection
// rules follow the reference converter (tutorials/ann/prepare_dif.c, file_dif.c:36-478):
//   * files whose first line mentions R060187 or "5.000", files without a peak table and files
//     measured at lambda = 0.710730 are skipped;
//   * "Sample ... T = x" gives the temperature (Celsius unless followed by K), default 25 C;
//   * the space group is the Hermann-Mauguin token after "SPACE GROUP:" looked up in a table
//     of 515 spellings (tools/spacegroups.inc).
// Cell parameters and atom positions of the DIF file are not needed for the sample and are not
// parsed here (the reference parses and then ignores them).
#include <dirent.h>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "spacegroups.inc"

static const double kMinTheta = 5.0, kMaxTheta = 90.0;

struct Dif {
  double temp = 273.15 + 25.0, lambda = 1.541838;
  int space = 0, n_peaks = 0;
  std::vector<double> t, i;  // raw pattern
};

static std::unordered_map<std::string, int> sg_table() {
  std::unordered_map<std::string, int> m;
  for (const auto &g : kSpaceGroups) {
    std::string s = g.symbols;
    size_t p = 0;
    while (p < s.size()) {
      size_t e = s.find(' ', p);
      if (e == std::string::npos) e = s.size();
      m[s.substr(p, e - p)] = g.number;
      p = e + 1;
    }
  }
  return m;
}

static bool getline_s(FILE *f, std::string &out) {
  static char *buf = nullptr;
  static size_t cap = 0;
  ssize_t n = getline(&buf, &cap, f);
  if (n < 0) return false;
  out.assign(buf, (size_t)n);
  return true;
}

static bool read_dif(const std::string &path, const std::unordered_map<std::string, int> &sg, Dif &d) {
  FILE *f = fopen(path.c_str(), "r+");  // "+" refuses directories, as in the reference
  if (!f) {
    fprintf(stderr, "Error opening file: %s\n", path.c_str());
    return false;
  }
  std::string line;
  bool ok = getline_s(f, line) && line.find("R060187") == std::string::npos && line.find("5.000") == std::string::npos;
  bool pending = false;  // `line` already holds the next unread line
  while (ok && (pending || getline_s(f, line))) {
    pending = false;
    const char *s = line.c_str();
    const char *p;
    if (strstr(s, "Sample") && (p = strstr(s, "T ="))) {
      char *end = nullptr;
      d.temp = strtod(p + 4, &end);
      if (!end || end[0] == '\0' || end[1] != 'K') d.temp += 273.15;
    }
    if ((p = strstr(s, "SPACE GROUP"))) {
      p = strchr(p, ':');
      if (p) {
        ++p;
        while (*p && !isgraph((unsigned char)*p)) ++p;
        const char *e = p;
        while (isgraph((unsigned char)*e)) ++e;
        auto it = sg.find(std::string(p, e));
        if (it != sg.end()) d.space = it->second;
        else fprintf(stdout, "#DBG: NO_space group = %s", p);
      }
    }
    if ((p = strstr(s, "WAVELENGTH"))) {
      while (*p && !isdigit((unsigned char)*p)) ++p;
      if (isdigit((unsigned char)*p)) d.lambda = strtod(p, nullptr);
    }
    if (strstr(s, "2-THETA")) {
      bool first = true;
      while (getline_s(f, line)) {
        const char *q = line.c_str();
        while (*q && !isgraph((unsigned char)*q) && *q != '\n') ++q;
        if (first)
          while (*q && !isdigit((unsigned char)*q) && *q != '\n') ++q;
        first = false;
        if (!isdigit((unsigned char)*q)) {
          pending = true;  // not a peak line: hand it back to the keyword scan
          break;
        }
        ++d.n_peaks;
      }
    }
  }
  fclose(f);
  return ok && d.n_peaks > 0;
}

static bool read_raw(const std::string &path, Dif &d) {
  FILE *f = fopen(path.c_str(), "r+");
  if (!f) {
    fprintf(stderr, "Error opening file: %s\n", path.c_str());
    return false;
  }
  std::string line;
  bool have = false;
  while (getline_s(f, line))
    if (isdigit((unsigned char)line[0])) {
      have = true;
      break;
    }
  while (have) {
    const char *s = line.c_str();
    char *e1 = nullptr, *e2 = nullptr;
    const double t = strtod(s, &e1);
    if (e1 != s && *e1) {
      const char *q = e1 + 1;
      while (*q && !isdigit((unsigned char)*q) && *q != '\n') ++q;  // a sign is skipped, as in the reference
      const double v = strtod(q, &e2);
      if (e2 != q) {
        d.t.push_back(t);
        d.i.push_back(v);
      }
    }
    have = getline_s(f, line);
  }
  fclose(f);
  return !d.t.empty();
}

static bool write_sample(const Dif &d, FILE *out, unsigned n_inputs, unsigned n_outputs) {
  if (!out || n_inputs < 2 || n_outputs == 0) return false;
  std::vector<double> bins(n_inputs - 1, 0.0);
  const double step = (kMaxTheta - kMinTheta) / (n_inputs - 1);
  double edge = kMinTheta + step, top = 0.0;
  size_t j = 0;
  while (j < d.t.size() && d.t[j] < kMinTheta) ++j;
  for (unsigned b = 0; b + 1 < n_inputs; ++b) {
    double acc = 0.0;
    while (j < d.t.size() && d.t[j] < edge) acc += d.i[j++];
    edge += step;
    bins[b] = acc;
    if (acc > top) top = acc;
  }
  if (top == 0.0) return false;
  fprintf(out, "[input] %u\n%7.5f", n_inputs, d.temp / 273.15);
  for (double v : bins) fprintf(out, " %7.5f", v / top);
  fprintf(out, "\n[output] %u\n", n_outputs);
  for (unsigned c = 0; c < n_outputs; ++c) fprintf(out, "%s%s", c ? " " : "", (int)c == d.space - 1 ? "1.0" : "-1.0");
  fprintf(out, "\n");
  return true;
}

static void usage() {
  fprintf(stdout, "usage: pdif rruff_dir -i N_BINS -o N_OUTPUTS [-s samples_dir]\n"
                  "rruff_dir must contain dif/ and raw/ with identically named files;\n"
                  "the sample gets N_BINS+1 inputs (temperature first).\n");
}

int main(int argc, char **argv) {
  unsigned n_in = 0, n_out = 0;
  std::string rruff, samples = "./samples";
  for (int a = 1; a < argc; ++a) {
    const char *s = argv[a];
    if (s[0] == '-' && s[1]) {
      const char opt = s[1];
      if (opt == 'h') return usage(), 0;
      const char *val = s[2] ? s + 2 : (a + 1 < argc ? argv[++a] : "");
      if (opt == 'i' || opt == 'o') {
        if (!isdigit((unsigned char)*val) || atoi(val) == 0) {
          fprintf(stderr, "syntax error: bad -%c parameter!\n", opt);
          return usage(), 1;
        }
        (opt == 'i' ? n_in : n_out) = (unsigned)atoi(val);
      } else if (opt == 's') {
        samples = val;
      } else {
        fprintf(stderr, "syntax error: unrecognized option!\n");
        return usage(), 1;
      }
    } else if (rruff.empty()) {
      rruff = s;
    } else {
      fprintf(stderr, "syntax error: too many parameters!\n");
      return usage(), 1;
    }
  }
  if (rruff.empty() || !n_in || !n_out) return usage(), 1;
  n_in += 1;  // temperature
  fprintf(stdout, ">> received: %s -i %u -o %u -s %s\n", rruff.c_str(), n_in, n_out, samples.c_str());
  DIR *probe = opendir(samples.c_str());
  if (!probe) {
    fprintf(stderr, "ERROR: can't open directory: %s\n", samples.c_str());
    return 1;
  }
  closedir(probe);
  const std::string dif_dir = rruff + "/dif/";
  DIR *dir = opendir(dif_dir.c_str());
  if (!dir) {
    fprintf(stderr, "ERROR: can't open directory: %s\n", dif_dir.c_str());
    return 1;
  }
  const auto sg = sg_table();
  while (struct dirent *e = readdir(dir)) {
    const std::string name = e->d_name;
    if (name == "." || name == "..") continue;
    fprintf(stdout, "Processing file: %s\n", name.c_str());
    Dif d;
    if (!read_dif(dif_dir + name, sg, d)) {
      fprintf(stderr, "ERROR:  reading %s file! SKIP\n", name.c_str());
      continue;
    }
    if (d.lambda == 0.710730) {
      fprintf(stderr, "ERROR:  file %s has wavelength of 0.710730! SKIP\n", name.c_str());
      continue;
    }
    const std::string raw = rruff + "/raw/" + name;
    if (!read_raw(raw, d)) {
      fprintf(stderr, "ERROR: reading %s file! SKIP\n", raw.c_str());
      continue;
    }
    const std::string dst = samples + "/" + name;
    FILE *out = fopen(dst.c_str(), "w");
    if (!out) {
      fprintf(stderr, "ERROR: opening %s sample file for WRITE!\n", dst.c_str());
      continue;
    }
    if (!write_sample(d, out, n_in, n_out)) fprintf(stderr, "ERROR: writting %s sample file!\n", dst.c_str());
    fclose(out);
  }
  closedir(dir);
  return 0;
}
