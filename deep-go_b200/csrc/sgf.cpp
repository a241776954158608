// SGF front end of makedata.lua: game text -> handicap stones, move list, dan ranks.  Host only, reentrant.
// Restates split_sgf (makedata.lua:40-57), to_move (:59-66), all_moves (:69-90), handicaps (:24-38), to_rank / get_ranks (:92-120)
// INCLUDING what their string handling implies, because the corpus the reference trained on is whatever these functions let through:
//   * the text is cut into lines at CR LF only (the pattern is the two-character sequence), then into pieces at ';';
//   * a piece counts as a (name, value) token only if it contains exactly one '[' and ends with ']' -- "GM[1]FF[4]" is dropped,
//     and so is a move followed by a bare line feed or by the closing ')' of the game;
//   * names are compared exactly ("B", "W", "BR", "WR"): no trimming;
//   * a move value that is not two characters, or whose first character is not a..s, is a pass and is skipped;
//   * handicap stones are read only from LINES that start with "AB" or "AW";
//   * a rank is a number followed by 'd'; a game without two such ranks yields no records (transcribe_from_to :548-556).
// Parity: unpinned by the reference (it ships no SGF file); the replay it feeds IS pinned (tests/test_makedata.py rebuilds the
// fixture games' records from SGF text synthesised out of the fixtures' own moves).
#include "../../include/deepgo.h"
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

void dg_set_error(const char *fmt, ...);

namespace {
typedef std::pair<const char *, size_t> Span;

// string:split(pat) for a literal separator: every segment between matches, empty ones included, the remainder last
std::vector<Span> split(Span s, const char *sep, size_t seplen) {
    std::vector<Span> out;
    const char *p = s.first, *end = s.first + s.second;
    for (;;) {
        const char *q = p;
        const char *hit = nullptr;
        for (; q + seplen <= end; ++q)
            if (memcmp(q, sep, seplen) == 0) { hit = q; break; }
        if (!hit) { out.push_back(Span(p, (size_t)(end - p))); return out; }
        out.push_back(Span(p, (size_t)(hit - p)));
        p = hit + seplen;
    }
}

struct Token { Span name, value; };

std::vector<Token> split_sgf(Span text) {
    std::vector<Token> out;
    for (const Span &line : split(text, "\r\n", 2))
        for (const Span &piece : split(line, ";", 1)) {
            const std::vector<Span> sub = split(piece, "[", 1);
            if (sub.size() == 2 && sub[1].second >= 1 && sub[1].first[sub[1].second - 1] == ']')
                out.push_back(Token{sub[0], Span(sub[1].first, sub[1].second - 1)});
        }
    return out;
}

bool is(Span s, const char *lit) { return s.second == strlen(lit) && memcmp(s.first, lit, s.second) == 0; }
int coord(char c) { return c >= 'a' && c <= 's' ? c - 'a' + 1 : 0; }       // chars_to_ints, 0 = nil

// to_rank: "<number>d" -> number (Lua tonumber: optional blanks around a decimal or hexadecimal numeral)
bool to_rank(Span s, double *out) {
    if (s.second < 1 || s.first[s.second - 1] != 'd') return false;
    std::string t(s.first, s.second - 1);
    const char *b = t.c_str();
    char *e = nullptr;
    const double v = strtod(b, &e);
    if (e == b) return false;
    while (*e == ' ' || *e == '\t' || *e == '\n' || *e == '\r' || *e == '\f' || *e == '\v') ++e;
    if (*e != 0) return false;
    *out = v;
    return true;
}
}  // namespace

extern "C" int dg_sgf_parse(const char *sgf, size_t len, int ranks[2], int *has_ranks, uint8_t *hx, uint8_t *hy, uint8_t *hplayer,
                            int max_handicaps, int *n_handicaps, uint8_t *mx, uint8_t *my, uint8_t *mplayer, int max_moves, int *n_moves) {
    if (!sgf || !ranks || !has_ranks || !n_handicaps || !n_moves || max_handicaps < 0 || max_moves < 0) { dg_set_error("dg_sgf_parse: bad argument"); return 1; }
    const Span text(sgf, len);
    const std::vector<Token> toks = split_sgf(text);
    // get_ranks: the last BR / WR wins; both must be dan ranks
    bool have[2] = {false, false};
    double rk[2] = {0, 0};
    for (const Token &t : toks) {
        const int who = is(t.name, "BR") ? 0 : (is(t.name, "WR") ? 1 : -1);
        if (who >= 0) have[who] = to_rank(t.value, &rk[who]);
    }
    *has_ranks = have[0] && have[1];
    ranks[0] = ranks[1] = 0;
    if (*has_ranks) {
        for (int i = 0; i < 2; ++i) {
            if (rk[i] != (double)(int)rk[i] || rk[i] < 0 || rk[i] > 255) { dg_set_error("dg_sgf_parse: rank %g is not a small integer", rk[i]); return 5; }
            ranks[i] = (int)rk[i];
        }
    }
    // handicaps: lines that start with AB / AW; positions = line:sub(4, -2) cut at "]["
    int nh = 0;
    for (const Span &line : split(text, "\r\n", 2)) {
        if (line.second < 2 || line.first[0] != 'A' || (line.first[1] != 'B' && line.first[1] != 'W')) continue;
        const int player = line.first[1] == 'W' ? 2 : 1;
        const Span body(line.second >= 4 ? line.first + 3 : line.first + line.second, line.second >= 4 ? line.second - 4 : 0);
        for (const Span &pos : split(body, "][", 2)) {
            const int x = pos.second == 2 ? coord(pos.first[0]) : 0, y = pos.second == 2 ? coord(pos.first[1]) : 0;
            if (!x || !y) { dg_set_error("dg_sgf_parse: malformed handicap position (the reference fails on it too)"); return 5; }
            if (nh >= max_handicaps) { dg_set_error("dg_sgf_parse: more than %d handicap stones", max_handicaps); return 2; }
            if (hx) { hx[nh] = (uint8_t)x; hy[nh] = (uint8_t)y; hplayer[nh] = (uint8_t)player; }
            ++nh;
        }
    }
    *n_handicaps = nh;
    // all_moves: B / W tokens; a value that does not start a coordinate is a pass
    int nm = 0;
    for (const Token &t : toks) {
        const int player = is(t.name, "B") ? 1 : (is(t.name, "W") ? 2 : 0);
        if (!player) continue;
        if (t.value.second != 2) continue;
        const int x = coord(t.value.first[0]), y = coord(t.value.first[1]);
        if (!x) continue;
        if (!y) { dg_set_error("dg_sgf_parse: move %d has a column but no row (the reference fails on it too)", nm + 1); return 5; }
        if (nm >= max_moves) { dg_set_error("dg_sgf_parse: more than %d moves", max_moves); return 2; }
        if (mx) { mx[nm] = (uint8_t)x; my[nm] = (uint8_t)y; mplayer[nm] = (uint8_t)player; }
        ++nm;
    }
    *n_moves = nm;
    return 0;
}
