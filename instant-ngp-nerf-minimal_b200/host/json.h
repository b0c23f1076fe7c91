// Minimal JSON value with a comment-tolerant parser and msgpack (de)serialisation.
// Stands in for nlohmann::json in the reference's host code (config files: internal/testbed.cu:44-92; snapshots:
// internal/testbed.cu:127-163, tiny-cuda-nn/gpu_memory_json.h:36-58 -- device buffers travel as msgpack "bin").
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace ngp {

class Json {
public:
	enum Type { Null, Bool, Int, Float, String, Binary, Array, Object };
	Type type = Null;
	bool b = false;
	int64_t i = 0;
	double f = 0.0;
	std::string s;                       // String, or raw bytes for Binary
	std::vector<Json> arr;
	std::vector<std::pair<std::string, Json>> obj;   // insertion-ordered

	Json() {}
	Json(bool v) : type(Bool), b(v) {}
	Json(int v) : type(Int), i(v) {}
	Json(uint32_t v) : type(Int), i(v) {}
	Json(int64_t v) : type(Int), i(v) {}
	Json(uint64_t v) : type(Int), i((int64_t)v) {}
	Json(double v) : type(Float), f(v) {}
	Json(float v) : type(Float), f(v) {}
	Json(const char* v) : type(String), s(v) {}
	Json(const std::string& v) : type(String), s(v) {}
	static Json binary(const void* data, size_t n) { Json j; j.type = Binary; j.s.assign((const char*)data, n); return j; }
	static Json array() { Json j; j.type = Array; return j; }
	static Json object() { Json j; j.type = Object; return j; }

	bool is_null() const { return type == Null; }
	bool is_object() const { return type == Object; }
	bool is_array() const { return type == Array; }
	bool is_number() const { return type == Int || type == Float; }
	size_t size() const { return type == Array ? arr.size() : type == Object ? obj.size() : 0; }

	bool contains(const std::string& k) const {
		if (type != Object) return false;
		for (auto& kv : obj) if (kv.first == k) return true;
		return false;
	}
	const Json& at(const std::string& k) const {
		if (type == Object) for (auto& kv : obj) if (kv.first == k) return kv.second;
		throw std::runtime_error("json: missing key '" + k + "'");
	}
	// like nlohmann: operator[] on a null/object creates the key
	Json& operator[](const std::string& k) {
		if (type == Null) type = Object;
		if (type != Object) throw std::runtime_error("json: not an object (key '" + k + "')");
		for (auto& kv : obj) if (kv.first == k) return kv.second;
		obj.emplace_back(k, Json());
		return obj.back().second;
	}
	const Json& operator[](const std::string& k) const { return at(k); }
	Json& operator[](size_t idx) { if (type != Array || idx >= arr.size()) throw std::runtime_error("json: bad array index"); return arr[idx]; }
	const Json& operator[](size_t idx) const { if (type != Array || idx >= arr.size()) throw std::runtime_error("json: bad array index"); return arr[idx]; }
	void push_back(const Json& v) { if (type == Null) type = Array; arr.push_back(v); }

	double number() const {
		if (type == Int) return (double)i;
		if (type == Float) return f;
		if (type == Bool) return b ? 1.0 : 0.0;
		throw std::runtime_error("json: not a number");
	}
	const std::string& str() const { if (type != String) throw std::runtime_error("json: not a string"); return s; }
	bool boolean() const { if (type == Bool) return b; if (is_number()) return number() != 0; throw std::runtime_error("json: not a bool"); }

	// config.value(key, default) of the reference
	float value(const std::string& k, float d) const { return contains(k) ? (float)at(k).number() : d; }
	double value(const std::string& k, double d) const { return contains(k) ? at(k).number() : d; }
	int value(const std::string& k, int d) const { return contains(k) ? (int)at(k).number() : d; }
	uint32_t value(const std::string& k, uint32_t d) const { return contains(k) ? (uint32_t)at(k).number() : d; }
	bool value(const std::string& k, bool d) const { return contains(k) ? at(k).boolean() : d; }
	std::string value(const std::string& k, const char* d) const { return contains(k) ? at(k).str() : std::string(d); }

	// ---- text ------------------------------------------------------------------------------------------
	static Json parse(const std::string& text) {
		Parser p{text.data(), text.data() + text.size()};
		Json j = p.value();
		p.ws();
		if (p.cur != p.end) p.fail("trailing characters");
		return j;
	}

	// ---- msgpack ---------------------------------------------------------------------------------------
	void to_msgpack(std::string& out) const {
		switch (type) {
			case Null: out.push_back((char)0xc0); break;
			case Bool: out.push_back((char)(b ? 0xc3 : 0xc2)); break;
			case Int:
				if (i >= 0) {
					if (i < 128) out.push_back((char)i);
					else if (i <= 0xff) { out.push_back((char)0xcc); be(out, (uint64_t)i, 1); }
					else if (i <= 0xffff) { out.push_back((char)0xcd); be(out, (uint64_t)i, 2); }
					else if (i <= 0xffffffffll) { out.push_back((char)0xce); be(out, (uint64_t)i, 4); }
					else { out.push_back((char)0xcf); be(out, (uint64_t)i, 8); }
				} else {
					if (i >= -32) out.push_back((char)i);
					else if (i >= -128) { out.push_back((char)0xd0); be(out, (uint64_t)i, 1); }
					else if (i >= -32768) { out.push_back((char)0xd1); be(out, (uint64_t)i, 2); }
					else if (i >= -2147483648ll) { out.push_back((char)0xd2); be(out, (uint64_t)i, 4); }
					else { out.push_back((char)0xd3); be(out, (uint64_t)i, 8); }
				}
				break;
			case Float: { out.push_back((char)0xcb); uint64_t u; std::memcpy(&u, &f, 8); be(out, u, 8); break; }
			case String:
				if (s.size() < 32) out.push_back((char)(0xa0 | s.size()));
				else if (s.size() <= 0xff) { out.push_back((char)0xd9); be(out, s.size(), 1); }
				else if (s.size() <= 0xffff) { out.push_back((char)0xda); be(out, s.size(), 2); }
				else { out.push_back((char)0xdb); be(out, s.size(), 4); }
				out += s; break;
			case Binary:
				if (s.size() <= 0xff) { out.push_back((char)0xc4); be(out, s.size(), 1); }
				else if (s.size() <= 0xffff) { out.push_back((char)0xc5); be(out, s.size(), 2); }
				else { out.push_back((char)0xc6); be(out, s.size(), 4); }
				out += s; break;
			case Array:
				if (arr.size() < 16) out.push_back((char)(0x90 | arr.size()));
				else if (arr.size() <= 0xffff) { out.push_back((char)0xdc); be(out, arr.size(), 2); }
				else { out.push_back((char)0xdd); be(out, arr.size(), 4); }
				for (auto& v : arr) v.to_msgpack(out);
				break;
			case Object:
				if (obj.size() < 16) out.push_back((char)(0x80 | obj.size()));
				else if (obj.size() <= 0xffff) { out.push_back((char)0xde); be(out, obj.size(), 2); }
				else { out.push_back((char)0xdf); be(out, obj.size(), 4); }
				for (auto& kv : obj) { Json(kv.first).to_msgpack(out); kv.second.to_msgpack(out); }
				break;
		}
	}
	static Json from_msgpack(const std::string& data) {
		size_t pos = 0;
		Json j = unpack(data, pos);
		return j;
	}

private:
	static void be(std::string& out, uint64_t v, int n) { for (int k = n - 1; k >= 0; --k) out.push_back((char)((v >> (8 * k)) & 0xff)); }
	static uint64_t rd(const std::string& d, size_t& pos, int n) {
		if (pos + n > d.size()) throw std::runtime_error("msgpack: truncated");
		uint64_t v = 0;
		for (int k = 0; k < n; ++k) v = (v << 8) | (uint8_t)d[pos++];
		return v;
	}
	static std::string bytes(const std::string& d, size_t& pos, size_t n) {
		if (pos + n > d.size()) throw std::runtime_error("msgpack: truncated");
		std::string r = d.substr(pos, n); pos += n; return r;
	}
	static Json unpack(const std::string& d, size_t& pos) {
		const uint8_t c = (uint8_t)rd(d, pos, 1);
		if (c < 0x80) return Json((int64_t)c);
		if (c >= 0xe0) return Json((int64_t)(int8_t)c);
		if ((c & 0xf0) == 0x80) return unpack_map(d, pos, c & 0x0f);
		if ((c & 0xf0) == 0x90) return unpack_arr(d, pos, c & 0x0f);
		if ((c & 0xe0) == 0xa0) return Json(bytes(d, pos, c & 0x1f));
		switch (c) {
			case 0xc0: return Json();
			case 0xc2: return Json(false);
			case 0xc3: return Json(true);
			case 0xc4: { size_t n = rd(d, pos, 1); std::string s = bytes(d, pos, n); return binary(s.data(), s.size()); }
			case 0xc5: { size_t n = rd(d, pos, 2); std::string s = bytes(d, pos, n); return binary(s.data(), s.size()); }
			case 0xc6: { size_t n = rd(d, pos, 4); std::string s = bytes(d, pos, n); return binary(s.data(), s.size()); }
			case 0xca: { uint32_t u = (uint32_t)rd(d, pos, 4); float v; std::memcpy(&v, &u, 4); return Json((double)v); }
			case 0xcb: { uint64_t u = rd(d, pos, 8); double v; std::memcpy(&v, &u, 8); return Json(v); }
			case 0xcc: return Json((int64_t)rd(d, pos, 1));
			case 0xcd: return Json((int64_t)rd(d, pos, 2));
			case 0xce: return Json((int64_t)rd(d, pos, 4));
			case 0xcf: return Json((int64_t)rd(d, pos, 8));
			case 0xd0: return Json((int64_t)(int8_t)rd(d, pos, 1));
			case 0xd1: return Json((int64_t)(int16_t)rd(d, pos, 2));
			case 0xd2: return Json((int64_t)(int32_t)rd(d, pos, 4));
			case 0xd3: return Json((int64_t)rd(d, pos, 8));
			case 0xd9: { size_t n = rd(d, pos, 1); return Json(bytes(d, pos, n)); }
			case 0xda: { size_t n = rd(d, pos, 2); return Json(bytes(d, pos, n)); }
			case 0xdb: { size_t n = rd(d, pos, 4); return Json(bytes(d, pos, n)); }
			case 0xdc: return unpack_arr(d, pos, rd(d, pos, 2));
			case 0xdd: return unpack_arr(d, pos, rd(d, pos, 4));
			case 0xde: return unpack_map(d, pos, rd(d, pos, 2));
			case 0xdf: return unpack_map(d, pos, rd(d, pos, 4));
			default: throw std::runtime_error("msgpack: unsupported type byte");
		}
	}
	static Json unpack_arr(const std::string& d, size_t& pos, size_t n) { Json j = array(); j.arr.reserve(n); for (size_t k = 0; k < n; ++k) j.arr.push_back(unpack(d, pos)); return j; }
	static Json unpack_map(const std::string& d, size_t& pos, size_t n) {
		Json j = object();
		for (size_t k = 0; k < n; ++k) { Json key = unpack(d, pos); if (key.type != String) throw std::runtime_error("msgpack: non-string key"); j.obj.emplace_back(key.s, unpack(d, pos)); }
		return j;
	}

	struct Parser {
		const char* cur; const char* end;
		[[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string("json parse error: ") + what); }
		void ws() {
			for (;;) {
				while (cur < end && (*cur == ' ' || *cur == '\t' || *cur == '\n' || *cur == '\r')) ++cur;
				if (cur + 1 < end && cur[0] == '/' && cur[1] == '/') { while (cur < end && *cur != '\n') ++cur; continue; }
				if (cur + 1 < end && cur[0] == '/' && cur[1] == '*') { cur += 2; while (cur + 1 < end && !(cur[0] == '*' && cur[1] == '/')) ++cur; cur = cur + 2 <= end ? cur + 2 : end; continue; }
				break;
			}
		}
		Json value() {
			ws();
			if (cur >= end) fail("unexpected end");
			switch (*cur) {
				case '{': return object();
				case '[': return array();
				case '"': return Json(string());
				case 't': lit("true"); return Json(true);
				case 'f': lit("false"); return Json(false);
				case 'n': lit("null"); return Json();
				default: return number();
			}
		}
		void lit(const char* w) { size_t n = strlen(w); if ((size_t)(end - cur) < n || strncmp(cur, w, n) != 0) fail("bad literal"); cur += n; }
		Json number() {
			const char* st = cur;
			bool is_float = false;
			if (cur < end && (*cur == '-' || *cur == '+')) ++cur;
			while (cur < end && ((*cur >= '0' && *cur <= '9') || *cur == '.' || *cur == 'e' || *cur == 'E' || *cur == '-' || *cur == '+')) {
				if (*cur == '.' || *cur == 'e' || *cur == 'E') is_float = true;
				++cur;
			}
			if (cur == st) fail("bad number");
			std::string t(st, cur);
			try {
				if (is_float) return Json(std::stod(t));
				return Json((int64_t)std::stoll(t));
			} catch (...) { fail("bad number"); }
		}
		std::string string() {
			++cur;
			std::string r;
			while (cur < end && *cur != '"') {
				if (*cur == '\\') {
					if (++cur >= end) fail("bad escape");
					switch (*cur) {
						case 'n': r.push_back('\n'); break; case 't': r.push_back('\t'); break; case 'r': r.push_back('\r'); break;
						case 'b': r.push_back('\b'); break; case 'f': r.push_back('\f'); break;
						case 'u': {
							if (end - cur < 5) fail("bad \\u escape");
							unsigned cp = (unsigned)std::stoul(std::string(cur + 1, cur + 5), nullptr, 16); cur += 4;
							if (cp < 0x80) r.push_back((char)cp);
							else if (cp < 0x800) { r.push_back((char)(0xc0 | (cp >> 6))); r.push_back((char)(0x80 | (cp & 0x3f))); }
							else { r.push_back((char)(0xe0 | (cp >> 12))); r.push_back((char)(0x80 | ((cp >> 6) & 0x3f))); r.push_back((char)(0x80 | (cp & 0x3f))); }
							break;
						}
						default: r.push_back(*cur);
					}
					++cur;
				} else r.push_back(*cur++);
			}
			if (cur >= end) fail("unterminated string");
			++cur;
			return r;
		}
		Json array() {
			++cur; Json j = Json::array();
			ws();
			if (cur < end && *cur == ']') { ++cur; return j; }
			for (;;) {
				j.arr.push_back(value());
				ws();
				if (cur < end && *cur == ',') { ++cur; ws(); if (cur < end && *cur == ']') { ++cur; return j; } continue; }
				if (cur < end && *cur == ']') { ++cur; return j; }
				fail("expected , or ]");
			}
		}
		Json object() {
			++cur; Json j = Json::object();
			ws();
			if (cur < end && *cur == '}') { ++cur; return j; }
			for (;;) {
				ws();
				if (cur >= end || *cur != '"') fail("expected key");
				std::string k = string();
				ws();
				if (cur >= end || *cur != ':') fail("expected :");
				++cur;
				j.obj.emplace_back(k, value());
				ws();
				if (cur < end && *cur == ',') { ++cur; ws(); if (cur < end && *cur == '}') { ++cur; return j; } continue; }
				if (cur < end && *cur == '}') { ++cur; return j; }
				fail("expected , or }");
			}
		}
	};
};

} // namespace ngp
