// Wavefront OBJ / MTL ingest for cupt_scene_load_obj.
//
// The reference parses OBJ through a vendored third-party loader (tiny_obj_loader v0.9.13, reference
// tiny_obj_loader.cc) called from loadOBJ (modelLoader.h:132).  This is an independent parser that keeps the
// observable semantics of that loader for the statements a path-tracing scene uses, so that the arrays handed
// to the hot path are identical (tests/test_scene_ingest.py compares against oracle/_ref):
//   * number parsing by decimal-mantissa accumulation in double, then * 5^e and ldexp (tiny_obj_loader.cc:127-241)
//   * `v`, `f` (i, i/j, i//k, i/j/k, negative = relative), fan triangulation (tiny_obj_loader.cc:361-399)
//   * a new shape on every `g`, `o` and `usemtl`; vertices de-duplicated per shape by (v, vt, vn) triple in
//     first-use order (tiny_obj_loader.cc:305-339,341-411)
//   * `mtllib` read from <mtl_dir><name>; materials flushed on `newmtl` and at end of file, a nameless default
//     material appended when the file ends (tiny_obj_loader.cc:413-615); `Kd`, `Ke` are the fields loadOBJ reads
//
// ATTRIBUTION.  parse_number, fix_index and next_corner below follow tiny_obj_loader's tryParseDouble, fixIndex and
// parseTriple step by step (same accumulation order, same `ldexp(mant * pow(5, e), e)` assembly) because bit-identical
// floats and the same vertex numbering are the point; they are derived from that third-party code, which is
//     Copyright 2012-2015, Syoyo Fujita.  Licensed under the 2-clause BSD license:
//     Redistribution and use in source and binary forms, with or without modification, are permitted provided that
//     the following conditions are met: 1. Redistributions of source code must retain the above copyright notice,
//     this list of conditions and the following disclaimer.  2. Redistributions in binary form must reproduce the
//     above copyright notice, this list of conditions and the following disclaimer in the documentation and/or other
//     materials provided with the distribution.
//     THIS SOFTWARE IS PROVIDED BY THE COPYRIGHT HOLDERS AND CONTRIBUTORS "AS IS" AND ANY EXPRESS OR IMPLIED
//     WARRANTIES, INCLUDING, BUT NOT LIMITED TO, THE IMPLIED WARRANTIES OF MERCHANTABILITY AND FITNESS FOR A
//     PARTICULAR PURPOSE ARE DISCLAIMED.  IN NO EVENT SHALL THE COPYRIGHT OWNER OR CONTRIBUTORS BE LIABLE FOR ANY
//     DIRECT, INDIRECT, INCIDENTAL, SPECIAL, EXEMPLARY, OR CONSEQUENTIAL DAMAGES (INCLUDING, BUT NOT LIMITED TO,
//     PROCUREMENT OF SUBSTITUTE GOODS OR SERVICES; LOSS OF USE, DATA, OR PROFITS; OR BUSINESS INTERRUPTION) HOWEVER
//     CAUSED AND ON ANY THEORY OF LIABILITY, WHETHER IN CONTRACT, STRICT LIABILITY, OR TORT (INCLUDING NEGLIGENCE OR
//     OTHERWISE) ARISING IN ANY WAY OUT OF THE USE OF THIS SOFTWARE, EVEN IF ADVISED OF THE POSSIBILITY OF SUCH DAMAGE.
// The rest of this file is an independent parser.
#include "host_common.h"
#include <cctype>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>

namespace cupt {
namespace {

struct Corner {
	int v, vt, vn;
	bool operator<(const Corner& o) const
	{
		if (v != o.v) return v < o.v;
		if (vn != o.vn) return vn < o.vn;
		return vt < o.vt;
	}
};

bool blank(char c) { return c == ' ' || c == '\t'; }
bool eol(char c) { return c == '\r' || c == '\n' || c == '\0'; }

// decimal text -> double the way the vendored loader does it (not correctly rounded; kept on purpose)
bool parse_number(const char* s, const char* end, double* out)
{
	if (s >= end)
		return false;
	double mant = 0.0;
	int expo = 0;
	char sign = '+', esign = '+';
	const char* p = s;
	int nread = 0;
	bool more = false;
	if (*p == '+' || *p == '-')
		sign = *p++;
	else if (!isdigit((unsigned char)*p))
		return false;
	while ((more = (p != end)) && isdigit((unsigned char)*p)) {
		mant *= 10;
		mant += (int)(*p - '0');
		++p; ++nread;
	}
	if (nread == 0)
		return false;
	if (more) {
		bool go_exp = false;
		if (*p == '.') {
			++p;
			nread = 1;
			while ((more = (p != end)) && isdigit((unsigned char)*p)) {
				mant += (int)(*p - '0') * pow(10.0, -nread);
				++nread; ++p;
			}
			go_exp = more;
		} else if (*p == 'e' || *p == 'E') {
			go_exp = true;
		}
		if (go_exp && (*p == 'e' || *p == 'E')) {
			++p;
			if ((more = (p != end)) && (*p == '+' || *p == '-'))
				esign = *p++;
			else if (!isdigit((unsigned char)*p))
				return false;
			nread = 0;
			while ((more = (p != end)) && isdigit((unsigned char)*p)) {
				expo = expo * 10 + (int)(*p - '0');
				++p; ++nread;
			}
			if (esign != '+')
				expo = -expo;
			if (nread == 0)
				return false;
		}
	}
	*out = (sign == '+' ? 1 : -1) * ldexp(mant * pow(5.0, expo), expo);
	return true;
}

float next_float(const char*& tok)
{
	tok += strspn(tok, " \t");
	const char* end = tok + strcspn(tok, " \t\r");
	double val = 0.0;
	parse_number(tok, end, &val);
	tok = end;
	return (float)val;
}

int fix_index(int idx, int n)
{
	if (idx > 0) return idx - 1;
	if (idx == 0) return 0;
	return n + idx;
}

Corner next_corner(const char*& tok, int nv, int nvn, int nvt)
{
	Corner c = {-1, -1, -1};
	c.v = fix_index(atoi(tok), nv);
	tok += strcspn(tok, "/ \t\r");
	if (tok[0] != '/')
		return c;
	++tok;
	if (tok[0] == '/') {  // i//k
		++tok;
		c.vn = fix_index(atoi(tok), nvn);
		tok += strcspn(tok, "/ \t\r");
		return c;
	}
	c.vt = fix_index(atoi(tok), nvt);
	tok += strcspn(tok, "/ \t\r");
	if (tok[0] != '/')
		return c;
	++tok;
	c.vn = fix_index(atoi(tok), nvn);
	tok += strcspn(tok, "/ \t\r");
	return c;
}

std::string first_word(const char* tok)
{
	tok += strspn(tok, " \t\r\n");
	size_t n = strcspn(tok, " \t\r\n");
	return std::string(tok, n);
}

bool getline_trim(std::istream& in, std::string& line)
{
	if (in.peek() == -1)
		return false;
	std::getline(in, line);
	if (!line.empty() && line[line.size() - 1] == '\r')
		line.erase(line.size() - 1);
	return true;
}

void reset_material(ObjMaterial& m)
{
	m.name.clear();
	for (int i = 0; i < 3; ++i)
		m.diffuse[i] = m.emission[i] = 0.f;
}

void read_mtl(std::istream& in, std::vector<ObjMaterial>& materials, std::map<std::string, int>& by_name)
{
	ObjMaterial cur;
	reset_material(cur);
	std::string line;
	while (getline_trim(in, line)) {
		const char* tok = line.c_str();
		tok += strspn(tok, " \t");
		if (tok[0] == '\0' || tok[0] == '#')
			continue;
		if (strncmp(tok, "newmtl", 6) == 0 && blank(tok[6])) {
			if (!cur.name.empty()) {
				by_name.insert(std::make_pair(cur.name, (int)materials.size()));
				materials.push_back(cur);
			}
			reset_material(cur);
			cur.name = first_word(tok + 7);
			continue;
		}
		if (tok[0] == 'K' && tok[1] == 'd' && blank(tok[2])) {
			tok += 2;
			cur.diffuse[0] = next_float(tok); cur.diffuse[1] = next_float(tok); cur.diffuse[2] = next_float(tok);
			continue;
		}
		if (tok[0] == 'K' && tok[1] == 'e' && blank(tok[2])) {
			tok += 2;
			cur.emission[0] = next_float(tok); cur.emission[1] = next_float(tok); cur.emission[2] = next_float(tok);
			continue;
		}
	}
	by_name.insert(std::make_pair(cur.name, (int)materials.size()));
	materials.push_back(cur);
}

// one face group -> one shape; per-shape vertex table in first-use order
bool flush_group(std::vector<ObjShape>& shapes, const std::vector<float>& v,
                 std::vector<std::vector<Corner> >& group, int material, const std::string& name)
{
	if (group.empty())
		return false;
	ObjShape sh;
	std::map<Corner, uint32_t> seen;
	for (size_t f = 0; f < group.size(); ++f) {
		const std::vector<Corner>& face = group[f];
		for (size_t k = 2; k < face.size(); ++k) {
			const Corner tri[3] = {face[0], face[k - 1], face[k]};
			for (int c = 0; c < 3; ++c) {
				std::map<Corner, uint32_t>::iterator it = seen.find(tri[c]);
				uint32_t id;
				if (it != seen.end()) {
					id = it->second;
				} else {
					id = (uint32_t)(sh.positions.size() / 3);
					for (int a = 0; a < 3; ++a)
						sh.positions.push_back(v[(size_t)3 * tri[c].v + a]);
					seen[tri[c]] = id;
				}
				sh.indices.push_back(id);
			}
			sh.material_ids.push_back(material);
		}
	}
	sh.name = name;
	shapes.push_back(sh);
	group.clear();
	return true;
}

}  // namespace

int parse_obj(const char* path, const char* mtl_dir, std::vector<ObjShape>& shapes,
              std::vector<ObjMaterial>& materials, std::string& err)
{
	shapes.clear();
	std::ifstream in(path);
	if (!in) {
		err = std::string("Cannot open file [") + path + "]";
		return CUPT_ERR_IO;
	}
	std::vector<float> v;
	int nvn = 0, nvt = 0;
	std::vector<std::vector<Corner> > group;
	std::map<std::string, int> by_name;
	std::string name, line;
	int material = -1;
	size_t lineno = 0;
	while (getline_trim(in, line)) {
		++lineno;
		const char* tok = line.c_str();
		tok += strspn(tok, " \t");
		if (tok[0] == '\0' || tok[0] == '#')
			continue;
		if (tok[0] == 'v' && blank(tok[1])) {
			tok += 2;
			float x = next_float(tok), y = next_float(tok), z = next_float(tok);
			v.push_back(x); v.push_back(y); v.push_back(z);
			continue;
		}
		if (tok[0] == 'v' && tok[1] == 'n' && blank(tok[2])) { ++nvn; continue; }
		if (tok[0] == 'v' && tok[1] == 't' && blank(tok[2])) { ++nvt; continue; }
		if (tok[0] == 'f' && blank(tok[1])) {
			tok += 2;
			tok += strspn(tok, " \t");
			std::vector<Corner> face;
			while (!eol(tok[0])) {
				Corner c = next_corner(tok, (int)(v.size() / 3), nvn, nvt);
				if (c.v < 0 || (size_t)c.v >= v.size() / 3) {
					char buf[160];
					snprintf(buf, sizeof(buf), "%s:%zu: face references vertex %d of %zu", path, lineno, c.v + 1, v.size() / 3);
					err = buf;
					return CUPT_ERR_PARSE;
				}
				face.push_back(c);
				tok += strspn(tok, " \t\r");
			}
			group.push_back(face);
			continue;
		}
		if (strncmp(tok, "usemtl", 6) == 0 && blank(tok[6])) {
			std::string m = first_word(tok + 7);
			flush_group(shapes, v, group, material, name);
			std::map<std::string, int>::iterator it = by_name.find(m);
			material = (it != by_name.end()) ? it->second : -1;
			continue;
		}
		if (strncmp(tok, "mtllib", 6) == 0 && blank(tok[6])) {
			std::string file = std::string(mtl_dir ? mtl_dir : "") + first_word(tok + 7);
			std::ifstream mtl(file.c_str());
			read_mtl(mtl, materials, by_name);
			if (!mtl.is_open()) {  // the vendored loader returns this as the error string and stops
				err = "WARN: Material file [ " + file + " ] not found. Created a default material.";
				return CUPT_ERR_IO;
			}
			continue;
		}
		if (tok[0] == 'g' && blank(tok[1])) {
			flush_group(shapes, v, group, material, name);
			std::vector<std::string> names;
			while (!eol(tok[0])) {
				tok += strspn(tok, " \t");
				size_t e = strcspn(tok, " \t\r");
				names.push_back(std::string(tok, e));
				tok += e;
				tok += strspn(tok, " \t\r");
			}
			name = names.size() > 1 ? names[1] : "";
			continue;
		}
		if (tok[0] == 'o' && blank(tok[1])) {
			flush_group(shapes, v, group, material, name);
			name = first_word(tok + 2);
			continue;
		}
	}
	flush_group(shapes, v, group, material, name);
	return CUPT_OK;
}

}  // namespace cupt
