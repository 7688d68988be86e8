// drt_json.h -- minimal JSON DOM for the glTF chunk (numbers kept as double, like the reference's
// tinygltf/nlohmann path, so double->float narrowing of factors matches Scene.cu:68-69).
#pragma once
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace drt {

struct JsonValue {
	enum Type { Null, Bool, Number, String, Array, Object } type = Null;
	bool b = false;
	double num = 0.0;
	std::string str;
	std::vector<JsonValue> arr;
	std::vector<std::pair<std::string, JsonValue>> obj;

	const JsonValue* get(const char* key) const
	{
		if (type != Object) return nullptr;
		for (const auto& kv : obj)
			if (kv.first == key) return &kv.second;
		return nullptr;
	}
	bool has(const char* key) const { return get(key) != nullptr; }
	int get_int(const char* key, int dflt) const
	{
		const JsonValue* v = get(key);
		return (v && v->type == Number && v->num >= -2147483648.0 && v->num <= 2147483647.0) ? (int)v->num : dflt; // (a cast out of range is undefined)
	}
	// A JSON number meant as a byte offset / length / count: true and *out set when it is a finite integer-valued number in
	// [0, 2^53] (a cast of anything else to size_t is undefined behaviour); a missing key gives dflt.
	bool get_size(const char* key, size_t dflt, size_t* out) const
	{
		const double v = get_num(key, (double)dflt);
		if (!(v >= 0.0) || !(v <= 9007199254740992.0)) return false; // (also rejects NaN)
		*out = (size_t)v;
		return true;
	}
	double get_num(const char* key, double dflt) const
	{
		const JsonValue* v = get(key);
		return (v && v->type == Number) ? v->num : dflt;
	}
	std::string get_str(const char* key, const char* dflt = "") const
	{
		const JsonValue* v = get(key);
		return (v && v->type == String) ? v->str : std::string(dflt);
	}
	size_t size() const { return type == Array ? arr.size() : (type == Object ? obj.size() : 0); }
};

class JsonParser {
public:
	JsonParser(const char* s, size_t n) : p_(s), end_(s + n) {}
	bool parse(JsonValue& out, std::string& err)
	{
		if (!value(out, 0)) { err = err_.empty() ? "JSON syntax error" : err_; return false; }
		ws();
		return true;
	}

private:
	const char* p_;
	const char* end_;
	std::string err_;

	void ws() { while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_; }
	bool fail(const char* m) { if (err_.empty()) err_ = m; return false; }

	bool value(JsonValue& v, int depth)
	{
		if (depth > 256) return fail("JSON nesting too deep");
		ws();
		if (p_ >= end_) return fail("unexpected end of JSON");
		char c = *p_;
		if (c == '{') return object(v, depth);
		if (c == '[') return array(v, depth);
		if (c == '"') { v.type = JsonValue::String; return string(v.str); }
		if (c == 't' && end_ - p_ >= 4 && !strncmp(p_, "true", 4)) { p_ += 4; v.type = JsonValue::Bool; v.b = true; return true; }
		if (c == 'f' && end_ - p_ >= 5 && !strncmp(p_, "false", 5)) { p_ += 5; v.type = JsonValue::Bool; v.b = false; return true; }
		if (c == 'n' && end_ - p_ >= 4 && !strncmp(p_, "null", 4)) { p_ += 4; v.type = JsonValue::Null; return true; }
		return number(v);
	}
	bool number(JsonValue& v)
	{
		const char* s = p_;
		while (p_ < end_ && (strchr("+-0123456789.eE", *p_) != nullptr)) ++p_;
		if (p_ == s) return fail("bad JSON token");
		std::string tmp(s, p_ - s);
		char* e = nullptr;
		v.num = strtod(tmp.c_str(), &e);
		if (!e || *e) return fail("bad JSON number");
		v.type = JsonValue::Number;
		return true;
	}
	static void utf8(std::string& out, unsigned cp)
	{
		if (cp < 0x80) out += (char)cp;
		else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
		else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
		else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
	}
	bool hex4(unsigned& v)
	{
		if (end_ - p_ < 4) return false;
		v = 0;
		for (int i = 0; i < 4; i++) {
			char c = *p_++;
			v <<= 4;
			if (c >= '0' && c <= '9') v |= c - '0';
			else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
			else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
			else return false;
		}
		return true;
	}
	bool string(std::string& out)
	{
		++p_; // opening quote
		out.clear();
		while (p_ < end_) {
			char c = *p_++;
			if (c == '"') return true;
			if (c != '\\') { out += c; continue; }
			if (p_ >= end_) break;
			char e = *p_++;
			switch (e) {
			case '"': out += '"'; break;
			case '\\': out += '\\'; break;
			case '/': out += '/'; break;
			case 'b': out += '\b'; break;
			case 'f': out += '\f'; break;
			case 'n': out += '\n'; break;
			case 'r': out += '\r'; break;
			case 't': out += '\t'; break;
			case 'u': {
				unsigned cp;
				if (!hex4(cp)) return fail("bad \\u escape");
				if (cp >= 0xD800 && cp < 0xDC00 && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
					p_ += 2;
					unsigned lo;
					if (!hex4(lo)) return fail("bad \\u escape");
					cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
				}
				utf8(out, cp);
				break;
			}
			default: return fail("bad escape");
			}
		}
		return fail("unterminated string");
	}
	bool array(JsonValue& v, int depth)
	{
		++p_;
		v.type = JsonValue::Array;
		ws();
		if (p_ < end_ && *p_ == ']') { ++p_; return true; }
		for (;;) {
			v.arr.emplace_back();
			if (!value(v.arr.back(), depth + 1)) return false;
			ws();
			if (p_ >= end_) return fail("unterminated array");
			if (*p_ == ',') { ++p_; continue; }
			if (*p_ == ']') { ++p_; return true; }
			return fail("expected , or ]");
		}
	}
	bool object(JsonValue& v, int depth)
	{
		++p_;
		v.type = JsonValue::Object;
		ws();
		if (p_ < end_ && *p_ == '}') { ++p_; return true; }
		for (;;) {
			ws();
			if (p_ >= end_ || *p_ != '"') return fail("expected object key");
			std::string key;
			if (!string(key)) return false;
			ws();
			if (p_ >= end_ || *p_ != ':') return fail("expected :");
			++p_;
			v.obj.emplace_back(key, JsonValue());
			if (!value(v.obj.back().second, depth + 1)) return false;
			ws();
			if (p_ >= end_) return fail("unterminated object");
			if (*p_ == ',') { ++p_; continue; }
			if (*p_ == '}') { ++p_; return true; }
			return fail("expected , or }");
		}
	}
};

} // namespace drt
