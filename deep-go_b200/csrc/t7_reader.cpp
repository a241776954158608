// Host helper: read one deep-go position record written by torch.save (Torch7 binary, little-endian) without
// Torch7.  Replaces torch.load in load_and_preprocess (dataloader.lua:123-125); record layout = flatten_data()
// output (dataloader.lua:30-39): { input = ByteTensor(9,19,19), move = {x, y, player}, ranks = {r1, r2}, flat = true }.
// Table key order on disk is Lua hash order, so the file is parsed generically (no fixed offsets).
// Reentrant: may be called from loader threads (data.lua:14-19 runs load_minibatch in a thread pool).
#include "../../include/deepgo.h"
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

void dg_set_error(const char *fmt, ...);

namespace {

struct Node {
    enum Kind { NIL, NUM, STR, BOOL, TABLE, TENSOR } kind = NIL;
    double num = 0;
    std::string str;
    std::map<std::string, int> fields;      // table: key (string, or number printed as integer) -> node index
    std::vector<int64_t> size, stride;
    int64_t offset = 0;
    std::vector<uint8_t> bytes;             // tensor: raw storage bytes
    int elem = 1;
    bool is_double = false;
};

struct Parser {
    const uint8_t *p, *end;
    std::vector<Node> nodes;
    std::map<int, int> memo;                // torch reference index -> node
    bool ok = true;

    bool need(size_t n) { if ((size_t)(end - p) < n) { ok = false; return false; } return true; }
    int32_t i32() { int32_t v = 0; if (need(4)) { memcpy(&v, p, 4); p += 4; } return v; }
    int64_t i64() { int64_t v = 0; if (need(8)) { memcpy(&v, p, 8); p += 8; } return v; }
    double f64() { double v = 0; if (need(8)) { memcpy(&v, p, 8); p += 8; } return v; }
    std::string str() { int32_t n = i32(); std::string s; if (n >= 0 && need((size_t)n)) { s.assign((const char *)p, n); p += n; } else ok = false; return s; }

    int parse(int depth = 0) {
        if (!ok || depth > 64) { ok = false; return -1; }
        const int tag = i32();
        Node n;
        switch (tag) {
        case 0: n.kind = Node::NIL; break;
        case 1: n.kind = Node::NUM; n.num = f64(); break;
        case 2: n.kind = Node::STR; n.str = str(); break;
        case 5: n.kind = Node::BOOL; n.num = i32(); break;
        case 3: {
            const int ref = i32();
            if (memo.count(ref)) return memo[ref];
            const int cnt = i32();
            n.kind = Node::TABLE;
            nodes.push_back(n);
            const int self = (int)nodes.size() - 1;
            memo[ref] = self;
            for (int i = 0; i < cnt && ok; ++i) {
                const int k = parse(depth + 1), v = parse(depth + 1);
                if (k < 0 || v < 0) { ok = false; break; }
                std::string key;
                if (nodes[k].kind == Node::STR) key = nodes[k].str;
                else if (nodes[k].kind == Node::NUM) key = std::to_string((long long)nodes[k].num);
                else continue;
                nodes[self].fields[key] = v;
            }
            return self;
        }
        case 4: {
            const int ref = i32();
            if (memo.count(ref)) return memo[ref];
            std::string ver = str();
            std::string cls = ver.rfind("V ", 0) == 0 ? str() : ver;
            if (cls.size() > 6 && cls.compare(cls.size() - 6, 6, "Tensor") == 0) {
                const int nd = i32();
                if (nd < 0 || nd > 8) { ok = false; return -1; }
                n.kind = Node::TENSOR;
                for (int i = 0; i < nd; ++i) n.size.push_back(i64());
                for (int i = 0; i < nd; ++i) n.stride.push_back(i64());
                n.offset = i64() - 1;
                nodes.push_back(n);
                const int self = (int)nodes.size() - 1;
                memo[ref] = self;
                const int st = parse(depth + 1);
                if (st >= 0 && nodes[st].kind == Node::TENSOR) { nodes[self].bytes = nodes[st].bytes; nodes[self].elem = nodes[st].elem; nodes[self].is_double = nodes[st].is_double; }
                return self;
            }
            if (cls.size() > 7 && cls.compare(cls.size() - 7, 7, "Storage") == 0) {
                int elem = 1;
                if (cls == "torch.ByteStorage" || cls == "torch.CharStorage") elem = 1;
                else if (cls == "torch.ShortStorage") elem = 2;
                else if (cls == "torch.IntStorage" || cls == "torch.FloatStorage" || cls == "torch.CudaStorage") elem = 4;   // cutorch writes floats
                else if (cls == "torch.LongStorage" || cls == "torch.DoubleStorage" || cls == "torch.CudaDoubleStorage") elem = 8;
                else { ok = false; return -1; }
                const int64_t cnt = i64();
                if (cnt < 0 || !need((size_t)cnt * elem)) { ok = false; return -1; }
                n.kind = Node::TENSOR;
                n.elem = elem;
                n.is_double = cls.find("Double") != std::string::npos;
                n.bytes.assign(p, p + cnt * elem);
                p += cnt * elem;
                nodes.push_back(n);
                memo[ref] = (int)nodes.size() - 1;
                return memo[ref];
            }
            // any other torch class (nn modules, criteria): the default serialisation writes the object's table next
            {
                nodes.push_back(n);
                const int self = (int)nodes.size() - 1;
                memo[ref] = self;
                const int body = parse(depth + 1);
                if (body >= 0 && nodes[body].kind == Node::TABLE) { nodes[self].kind = Node::TABLE; nodes[self].fields = nodes[body].fields; }
                return self;
            }
        }
        case 6: case 7: case 8: {
            // a Lua function (Experiment:new copies the prototype's methods into every instance, so a pickled experiment carries
            // them): [reference index -- tags 7 and 8 only; the older tag 6 has none,] string.dump bytes, then the upvalues.  Skipped.
            int ref = -1;
            if (tag != 6) {
                ref = i32();
                if (memo.count(ref)) return memo[ref];
            }
            const int32_t nbytes = i32();
            if (nbytes < 0 || !need((size_t)nbytes)) { ok = false; return -1; }
            p += nbytes;
            nodes.push_back(n);
            const int self = (int)nodes.size() - 1;
            if (tag != 6) memo[ref] = self;
            parse(depth + 1);
            return self;
        }
        default: ok = false; return -1;
        }
        nodes.push_back(n);
        return (int)nodes.size() - 1;
    }
};

}  // namespace

extern "C" int dg_t7_read_record(const char *path, uint8_t *raw, int *x, int *y, int *player, int ranks[2]) {
    if (!path || !raw || !x || !y || !player || !ranks) { dg_set_error("dg_t7_read_record: NULL argument"); return 2; }
    FILE *f = fopen(path, "rb");
    if (!f) { dg_set_error("dg_t7_read_record: cannot open %s", path); return 4; }
    std::vector<uint8_t> buf;
    uint8_t tmp[8192];
    size_t got;
    while ((got = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + got);
    fclose(f);
    Parser ps;
    ps.p = buf.data();
    ps.end = buf.data() + buf.size();
    const int root = ps.parse();
    if (!ps.ok || root < 0 || ps.nodes[root].kind != Node::TABLE) { dg_set_error("dg_t7_read_record: %s is not a Torch7 table record", path); return 5; }
    auto &flds = ps.nodes[root].fields;
    if (!flds.count("input") || !flds.count("move") || !flds.count("ranks")) { dg_set_error("dg_t7_read_record: %s lacks input/move/ranks", path); return 5; }
    const Node &in = ps.nodes[flds["input"]];
    if (in.kind != Node::TENSOR || in.elem != 1 || in.size.size() != 3 || in.size[0] != 9 || in.size[1] != 19 || in.size[2] != 19) {
        dg_set_error("dg_t7_read_record: %s: input is not a ByteTensor(9,19,19)", path);
        return 5;
    }
    for (int a = 0; a < 9; ++a)
        for (int b = 0; b < 19; ++b)
            for (int c = 0; c < 19; ++c) {
                const int64_t idx = in.offset + a * in.stride[0] + b * in.stride[1] + c * in.stride[2];
                if (idx < 0 || idx >= (int64_t)in.bytes.size()) { dg_set_error("dg_t7_read_record: %s: tensor out of range", path); return 5; }
                raw[(a * 19 + b) * 19 + c] = in.bytes[idx];
            }
    const Node &mv = ps.nodes[flds["move"]], &rk = ps.nodes[flds["ranks"]];
    auto num = [&](const Node &t, const char *k, int *out) -> bool {
        auto it = t.fields.find(k);
        if (t.kind != Node::TABLE || it == t.fields.end() || ps.nodes[it->second].kind != Node::NUM) return false;
        *out = (int)ps.nodes[it->second].num;
        return true;
    };
    if (!num(mv, "x", x) || !num(mv, "y", y) || !num(mv, "player", player) || !num(rk, "1", &ranks[0]) || !num(rk, "2", &ranks[1])) {
        dg_set_error("dg_t7_read_record: %s: move/ranks fields missing", path);
        return 5;
    }
    return 0;
}

// experiment:load of a checkpoint the REFERENCE wrote (experiments.lua:124-131: torch.save of the pickled experiment table): the
// flat parameter vector is the field modelParameters (model:getParameters(), experiments.lua:107; a Float, Cuda or Double tensor),
// the rate lives in optimizer.rate (optimizer.lua:18-21).  Values missing from the file are left untouched.
extern "C" int dg_t7_read_model(const char *path, float *params, size_t cap, size_t *n_out, double *rate, double *rate_decay,
                                long long *iterations) {
    if (!path || !n_out) { dg_set_error("dg_t7_read_model: NULL argument"); return 2; }
    FILE *f = fopen(path, "rb");
    if (!f) { dg_set_error("dg_t7_read_model: cannot open %s", path); return 4; }
    std::vector<uint8_t> buf;
    uint8_t tmp[1 << 16];
    size_t got;
    while ((got = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + got);
    fclose(f);
    Parser ps;
    ps.p = buf.data(); ps.end = buf.data() + buf.size();
    const int root = ps.parse();
    if (!ps.ok || root < 0 || ps.nodes[root].kind != Node::TABLE) { dg_set_error("dg_t7_read_model: %s is not a Torch7 table", path); return 5; }
    auto &flds = ps.nodes[root].fields;
    if (!flds.count("modelParameters") || ps.nodes[flds["modelParameters"]].kind != Node::TENSOR) {
        dg_set_error("dg_t7_read_model: %s has no modelParameters tensor", path);
        return 5;
    }
    const Node &t = ps.nodes[flds["modelParameters"]];
    if (t.size.size() != 1 || (t.elem != 4 && t.elem != 8) || t.stride[0] != 1 || t.offset < 0 ||
        (uint64_t)(t.offset + t.size[0]) * t.elem > t.bytes.size()) {
        dg_set_error("dg_t7_read_model: %s: modelParameters is not a contiguous 1-D float tensor", path);
        return 5;
    }
    *n_out = (size_t)t.size[0];
    if (params) {
        if (cap < *n_out) { dg_set_error("dg_t7_read_model: buffer holds %zu values, the file has %zu", cap, *n_out); return 2; }
        for (size_t i = 0; i < *n_out; ++i) {
            if (t.elem == 4 && !t.is_double) { float v; memcpy(&v, &t.bytes[(t.offset + i) * 4], 4); params[i] = v; }
            else { double v; memcpy(&v, &t.bytes[(t.offset + i) * 8], 8); params[i] = (float)v; }
        }
    }
    if (flds.count("optimizer") && ps.nodes[flds["optimizer"]].kind == Node::TABLE) {
        auto &of = ps.nodes[flds["optimizer"]].fields;
        if (rate && of.count("rate") && ps.nodes[of["rate"]].kind == Node::NUM) *rate = ps.nodes[of["rate"]].num;
        if (rate_decay && of.count("rate_decay") && ps.nodes[of["rate_decay"]].kind == Node::NUM) *rate_decay = ps.nodes[of["rate_decay"]].num;
    }
    if (iterations && flds.count("iterations") && ps.nodes[flds["iterations"]].kind == Node::NUM) *iterations = (long long)ps.nodes[flds["iterations"]].num;
    return 0;
}
