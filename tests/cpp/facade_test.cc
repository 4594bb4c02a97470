// Drives the C++ facade (wormhole_b200/host) the way the reference's own code drives its headers:
// Localizer::Localize, SpMV::Times / TransTimes, CreateLoss -> Init / Evaluate / CalcGrad,
// KVWorker::ZPull / ZPush against an AsgdServer, and AsgdWorker::ProcessMinibatch.
// Reads minibatches from a binary file written by tests/test_cpp_facade.py, writes every result to
// another one; the Python test compares them with the CPU oracle.
#define WH_DEFINE_FLAGS
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>

#include "base/binary_class_evaluation.h"
#include "base/localizer.h"
#include "base/spmm.h"
#include "base/spmv.h"
#include "linear/async_sgd.h"

using namespace dmlc;

template <typename T> static void Put(FILE* f, const std::vector<T>& v) {
  uint64_t n = v.size(); fwrite(&n, 8, 1, f); if (n) fwrite(v.data(), sizeof(T), n, f);
}
template <typename T> static std::vector<T> Get(FILE* f) {
  uint64_t n = 0; if (fread(&n, 8, 1, f) != 1) return {};
  std::vector<T> v(n); if (n && fread(v.data(), sizeof(T), n, f) != n) v.clear(); return v;
}

// facade_test --reader file format mb_size shuf_buf neg_sampling seed max_key out.bin:
// MinibatchIter over a text file; every minibatch is written out (host view) and trained on (device view).
static int ReaderMode(int argc, char** argv) {
  if (argc < 10) { fprintf(stderr, "usage: facade_test --reader file fmt mb shuf neg seed max_key out.bin\n"); return 2; }
  const char* file = argv[2]; const char* fmt = argv[3];
  const unsigned mb = (unsigned)atoi(argv[4]), shuf = (unsigned)atoi(argv[5]);
  const float neg = (float)atof(argv[6]);
  srand((unsigned)atoi(argv[7]));
  ps::FLAGS_max_key = strtoull(argv[8], NULL, 10);
  FILE* out = fopen(argv[9], "wb");
  linear::Config conf; std::string err;
  conf.ParseFromString("algo = FTRL\nlambda_l1 = .5\nlambda_l2 = .1\nlr_eta = .1\nlr_beta = 1\n", &err);
  linear::AsgdServer server(conf);
  linear::AsgdWorker worker(conf, &server);
  std::vector<double> progs;
  worker.reporter = [&](const linear::Progress& p) { progs.insert(progs.end(), p.data.begin(), p.data.end()); };
  data::MinibatchIter<uint64_t> reader(file, 0, 1, fmt, mb, shuf, neg, 40000);
  reader.BeforeFirst();
  uint64_t n = 0;
  while (reader.Next()) {
    const RowBlock<uint64_t>& b = reader.Value();
    const size_t nnz = b.offset[b.size];
    Put(out, std::vector<size_t>(b.offset, b.offset + b.size + 1));
    Put(out, std::vector<uint64_t>(b.index, b.index + nnz));
    Put(out, b.value ? std::vector<float>(b.value, b.value + nnz) : std::vector<float>());
    Put(out, std::vector<float>(b.label, b.label + b.size));
    worker.ProcessMinibatch(reader.DeviceValue(), true);
    ++n;
  }
  Put(out, std::vector<size_t>());   // end marker: an empty offset array
  Put(out, progs);
  std::vector<int64_t> tail = {server.nnz_w(), (int64_t)reader.BytesRead()};
  Put(out, tail);
  fclose(out);
  printf("facade_test reader ok: %llu minibatches\n", (unsigned long long)n);
  return 0;
}

int main(int argc, char** argv) {
  if (argc > 1 && !strcmp(argv[1], "--reader")) return ReaderMode(argc, argv);
  if (argc < 4) { fprintf(stderr, "usage: facade_test in.bin out.bin conf.txt\n"); return 2; }
  FILE* in = fopen(argv[1], "rb"); FILE* out = fopen(argv[2], "wb");
  // conf in the reference's text format
  std::string text; { FILE* c = fopen(argv[3], "r"); char buf[4096]; size_t k; while ((k = fread(buf, 1, sizeof(buf), c)) > 0) text.append(buf, k); fclose(c); }
  linear::Config conf; std::string err;
  if (!conf.ParseFromString(text, &err)) { fprintf(stderr, "conf: %s\n", err.c_str()); return 3; }
  std::vector<uint64_t> hdr = Get<uint64_t>(in);   // {max_key, number of minibatches}
  ps::FLAGS_max_key = hdr[0];
  const int nmb = (int)hdr[1];

  linear::AsgdServer server(conf);
  linear::AsgdWorker worker(conf, &server);
  std::vector<double> progs;
  worker.reporter = [&](const linear::Progress& p) { progs.insert(progs.end(), p.data.begin(), p.data.end()); };
  ps::KVWorker<float> kv(server.server());
  std::vector<uint64_t> all_keys;

  for (int m = 0; m < nmb; ++m) {
    std::vector<size_t> offset = Get<size_t>(in);
    std::vector<uint64_t> index = Get<uint64_t>(in);
    std::vector<float> value = Get<float>(in), label = Get<float>(in);
    RowBlock<uint64_t> mb;
    mb.size = offset.size() - 1; mb.offset = offset.data(); mb.index = index.data();
    mb.value = value.empty() ? NULL : value.data(); mb.label = label.data();

    if (m == 0) {
      // the primitives one by one, as AsgdWorker::ProcessMinibatch strings them together
      data::RowBlockContainer<unsigned> data;
      auto feaid = std::make_shared<std::vector<uint64_t>>();
      std::vector<unsigned> cnt;
      Localizer<uint64_t> lc(2);
      lc.Localize(mb, &data, feaid.get(), &cnt);
      Put(out, *feaid); Put(out, cnt); Put(out, data.offset); Put(out, data.index);
      std::vector<float> val;
      kv.ZPull(feaid, &val);                       // all zero: a fresh model
      for (size_t i = 0; i < val.size(); ++i) val[i] = 0.01f * (float)((i * 2654435761u) % 97) - 0.4f;
      RowBlock<unsigned> D = data.GetBlock();
      std::vector<float> Xw(D.size), g(val.size());
      SpMV::Times(D, val, &Xw);
      std::vector<float> d(D.size);
      for (size_t i = 0; i < d.size(); ++i) d[i] = 0.001f * (float)(i % 1000) - 0.5f;
      SpMV::TransTimes(D, d, &g);
      Put(out, Xw); Put(out, g);
      auto loss = linear::CreateLoss<float>(conf.loss());
      loss->Init(D, val, 2);
      linear::Progress prog; loss->Evaluate(&prog);
      std::vector<float> grad;
      loss->CalcGrad(&grad);
      delete loss;
      Put(out, prog.data); Put(out, grad);
      BinClassEval<float> ev(D.label, Xw.data(), Xw.size());
      std::vector<float> e = {ev.AUC(), ev.Accuracy(0), ev.LogitObjv()};
      Put(out, e);
      const int dim = 4;
      std::vector<float> V(val.size() * dim), XV(D.size * dim), GV(val.size() * dim);
      for (size_t i = 0; i < V.size(); ++i) V[i] = 0.001f * (float)(i % 613) - 0.3f;
      SpMM::Times(D, V, &XV);
      SpMM::TransTimes(D, XV, &GV);
      Put(out, XV); Put(out, GV);
    }
    worker.ProcessMinibatch(mb, true);
    // remember the transformed keys for the final model dump
    for (uint64_t id : index) all_keys.push_back(ps::FLAGS_max_key == ~0ULL ? ReverseBytes(id) : id % ps::FLAGS_max_key);
  }
  std::sort(all_keys.begin(), all_keys.end());
  all_keys.erase(std::unique(all_keys.begin(), all_keys.end()), all_keys.end());
  std::vector<float> w, z, cg;
  server.State(all_keys, &w, &z, &cg);
  std::vector<int64_t> nnz = {server.nnz_w()};
  Put(out, progs); Put(out, all_keys); Put(out, w); Put(out, z); Put(out, cg); Put(out, nnz);
  fclose(in); fclose(out);
  printf("facade_test ok: %d minibatches, |w|_0 = %lld\n", nmb, (long long)nnz[0]);
  return 0;
}
