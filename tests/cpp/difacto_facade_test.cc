// Drives the C++ DiFacto facade (wormhole_b200/host/difacto/async_sgd.h): AsyncSGD::ProcessMinibatch over the
// minibatches of a binary file written by tests/test_cpp_facade.py, two data passes; writes the progress vectors and
// two progress-table rows. The Python test compares them with the ctypes path on the same inputs.
#define WH_DEFINE_FLAGS
#include <cstdio>
#include <vector>

#include <algorithm>
#include <memory>

#include "base/localizer.h"
#include "difacto/async_sgd.h"
#include "difacto/loss.h"

using namespace dmlc;

template <typename T> static void Put(FILE* f, const std::vector<T>& v) {
  uint64_t n = v.size(); fwrite(&n, 8, 1, f); if (n) fwrite(v.data(), sizeof(T), n, f);
}
template <typename T> static std::vector<T> Get(FILE* f) {
  uint64_t n = 0; if (fread(&n, 8, 1, f) != 1) return {};
  std::vector<T> v(n); if (n && fread(v.data(), sizeof(T), n, f) != n) v.clear(); return v;
}

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: difacto_facade_test in.bin out.bin\n"); return 2; }
  FILE* in = fopen(argv[1], "rb"); FILE* out = fopen(argv[2], "wb");
  std::vector<uint64_t> hdr = Get<uint64_t>(in);   // {max_key, number of minibatches}
  ps::FLAGS_max_key = hdr[0];
  difacto::Config conf;
  conf.lambda_l1 = 1; conf.lr_eta = .01f; conf.embedding.dim = 5; conf.embedding.threshold = 5; conf.embedding.init_scale = .01f;
  difacto::AsyncSGD job(conf, /*seed=*/7);
  std::vector<double> progs;
  difacto::Progress table;
  std::vector<std::string> rows;
  job.reporter = [&](const difacto::Progress& p) {
    progs.insert(progs.end(), p.data.begin(), p.data.end());
    table.data = p.data;
    rows.push_back(table.PrintStr());
  };
  struct MB { std::vector<size_t> offset; std::vector<uint64_t> index; std::vector<float> value, label; };
  std::vector<MB> mbs(hdr[1]);
  for (auto& m : mbs) { m.offset = Get<size_t>(in); m.index = Get<uint64_t>(in); m.value = Get<float>(in); m.label = Get<float>(in); }
  for (int pass = 0; pass < 2; ++pass)
    for (auto& m : mbs) {
      RowBlock<uint64_t> b;
      b.size = m.offset.size() - 1; b.offset = m.offset.data(); b.index = m.index.data();
      b.value = m.value.empty() ? NULL : m.value.data(); b.label = m.label.data();
      job.ProcessMinibatch(b, true, pass);
    }
  // The same job again, this time the way the reference's worker talks to its servers (learn/difacto/async_sgd.h:
  // 363-425): Localize with counts, ZPush(cmd = kPushFeaCnt), ZVPull depending on it, Loss::Evaluate / CalcGrad, ZVPush
  // -- against a second shard with the same seed. Progress and final model must equal the one-call path's.
  difacto::AsyncSGD shard(conf, /*seed=*/7);
  ps::KVWorker<float> server;
  server.set_difacto(shard.handle(), conf.embedding.dim);
  std::vector<double> progs2;
  std::vector<uint64_t> all_keys;
  for (int pass = 0; pass < 2; ++pass)
    for (auto& m : mbs) {
      RowBlock<uint64_t> b;
      b.size = m.offset.size() - 1; b.offset = m.offset.data(); b.index = m.index.data();
      b.value = m.value.empty() ? NULL : m.value.data(); b.label = m.label.data();
      data::RowBlockContainer<unsigned> data;
      auto feaid = std::make_shared<std::vector<uint64_t>>();
      auto feacnt = std::make_shared<std::vector<float>>();
      Localizer<uint64_t> lc;
      lc.Localize(b, &data, feaid.get(), feacnt.get());
      ps::SyncOpts pull_w_opt;
      if (pass == 0) {
        ps::SyncOpts cnt_opt;
        cnt_opt.cmd = ps::KVWorker<float>::kPushFeaCnt;
        pull_w_opt.deps.push_back(server.ZPush(feaid, feacnt, cnt_opt));
      }
      auto val = std::make_shared<std::vector<float>>();
      auto val_siz = std::make_shared<std::vector<int>>();
      bool pulled = false;
      pull_w_opt.callback = [&]() { pulled = true; };
      const int ts = server.ZVPull(feaid, val.get(), val_siz.get(), pull_w_opt);
      server.Wait(ts);
      if (!pulled) { fprintf(stderr, "ZVPull did not run its callback\n"); return 1; }
      difacto::Loss<float> loss(data.GetBlock(), *val, *val_siz, conf);
      difacto::Progress prog;
      loss.Evaluate(&prog);
      loss.CalcGrad(val.get());
      server.ZVPush(feaid, val, val_siz);
      progs2.insert(progs2.end(), prog.data.begin(), prog.data.end());
      all_keys.insert(all_keys.end(), feaid->begin(), feaid->end());
    }
  std::sort(all_keys.begin(), all_keys.end());
  all_keys.erase(std::unique(all_keys.begin(), all_keys.end()), all_keys.end());
  const int dim = conf.embedding.dim;
  for (whDifacto* d : {job.handle(), shard.handle()}) {
    const size_t n = all_keys.size();
    std::vector<uint32_t> cnt(n); std::vector<int> size(n);
    std::vector<float> w(n * (dim + 1)), cg(n * (dim + 2));
    WH_CHECK(wh_difacto_get_state(d, all_keys.data(), n, cnt.data(), size.data(), w.data(), cg.data()));
    Put(out, cnt); Put(out, size); Put(out, w); Put(out, cg);
  }
  Put(out, progs2);
  Put(out, progs);
  std::string joined;
  for (auto& r : rows) joined += r + "\n";
  Put(out, std::vector<char>(joined.begin(), joined.end()));
  fclose(in); fclose(out);
  printf("difacto_facade_test ok: %zu minibatches\n", progs.size() / 8);
  return 0;
}
