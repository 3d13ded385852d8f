// pmnist: MNIST idx3/idx1 files -> one hpnn text sample per image.
//
//   pmnist [--scale] [--ref-quirks] samples_dir tests_dir
//
// reads ./train_images ./train_labels ./test_images ./test_labels (the renamed MNIST files),
// writes samples_dir/sNNNNN.txt and tests_dir/sNNNNN.txt with 784 inputs ("%7.5f") and a
// +-1 one-hot over 10 outputs, label as a "#d" comment -- the format of the reference's
// converter (tutorials/mnist/prepare_mnist.c:47-60).
// Differences, on purpose:
//   * the reference reads the first test label twice, shifting every test label by one
//     (prepare_mnist.c:228-230); fixed here, --ref-quirks restores it;
//   * pixels are written raw 0..255 like the reference unless --scale (-> [0,1]) is given.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

static uint32_t be32(FILE *f, bool &ok) {
  unsigned char b[4];
  if (fread(b, 1, 4, f) != 4) {
    ok = false;
    return 0;
  }
  return (uint32_t(b[0]) << 24) | (uint32_t(b[1]) << 16) | (uint32_t(b[2]) << 8) | uint32_t(b[3]);
}

static int convert(const char *labels, const char *images, const std::string &out_dir, int &index, bool scale,
                   bool double_first_label, const char *what) {
  FILE *fl = fopen(labels, "rb"), *fi = fopen(images, "rb");
  if (!fl || !fi) {
    fprintf(stderr, "FAILED to open %s file %s for READ!\n", fl ? "image" : "label", fl ? images : labels);
    return -1;
  }
  bool ok = true;
  const uint32_t lmagic = be32(fl, ok), lcount = be32(fl, ok);
  const uint32_t imagic = be32(fi, ok), icount = be32(fi, ok);
  const uint32_t rows = be32(fi, ok), cols = be32(fi, ok);
  if (!ok || lcount != icount || rows * cols == 0) {
    fprintf(stderr, "ERROR: inconsistent MNIST headers (%u labels, %u images, %ux%u)\n", lcount, icount, rows, cols);
    return -1;
  }
  fprintf(stdout, "# Opened %s label=%X image=%X\n", what, lmagic, imagic);
  std::vector<unsigned char> px(rows * cols);
  unsigned char label = 0;
  if (double_first_label && fread(&label, 1, 1, fl) != 1) return -1;
  int written = 0;
  while (fread(&label, 1, 1, fl) == 1 && fread(px.data(), 1, px.size(), fi) == px.size()) {
    ++index;
    if (label > 9) {
      fprintf(stderr, "ERROR: label out of boundaries!\n");
      continue;
    }
    char name[32];
    snprintf(name, sizeof(name), "/s%05d.txt", index);
    FILE *fo = fopen((out_dir + name).c_str(), "w");
    if (!fo) {
      fprintf(stderr, "FAILED to open sample %s%s for WRITE!\n", out_dir.c_str(), name);
      return -1;
    }
    fprintf(fo, "[input] %u\n", rows * cols);
    for (size_t i = 0; i < px.size(); ++i)
      fprintf(fo, i ? " %7.5f" : "%7.5f", scale ? px[i] / 255.0 : (double)px[i]);
    fprintf(fo, "\n[output] %i  #%d\n", 10, (int)label);
    for (int c = 0; c < 10; ++c) fprintf(fo, "%s%s", c ? " " : "", c == label ? "1.0" : "-1.0");
    fprintf(fo, "\n");
    fclose(fo);
    ++written;
  }
  fclose(fl);
  fclose(fi);
  return written;
}

int main(int argc, char **argv) {
  bool scale = false, quirks = false;
  std::vector<std::string> dirs;
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "--scale")) scale = true;
    else if (!strcmp(argv[i], "--ref-quirks")) quirks = true;
    else if (!strcmp(argv[i], "-h") || !strcmp(argv[i], "--help")) dirs.clear(), i = argc;
    else dirs.push_back(argv[i]);
  }
  if (dirs.size() != 2) {
    fprintf(stdout, "usage: pmnist [--scale] [--ref-quirks] samples_dir tests_dir\n"
                    "expects ./train_images ./train_labels ./test_images ./test_labels\n"
                    "(train-images-idx3-ubyte, train-labels-idx1-ubyte, t10k-images-idx3-ubyte,\n"
                    " t10k-labels-idx1-ubyte renamed)\n");
    return dirs.empty() ? 0 : 1;
  }
  fprintf(stdout, "processing sample database into %s directory.\n", dirs[0].c_str());
  fprintf(stdout, "processing   test database into %s directory.\n", dirs[1].c_str());
  int index = 0;  // the reference keeps counting across both sets: tests start at s60001.txt
  if (convert("./train_labels", "./train_images", dirs[0], index, scale, false, "samples") < 0) return -1;
  if (convert("./test_labels", "./test_images", dirs[1], index, scale, quirks, "tests") < 0) return -1;
  return 0;
}
