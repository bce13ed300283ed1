// osc_urdf.cpp -- URDF (+ world file) -> serial-chain constants on the host (C++, no dependency): SURVEY.md 8(f) row f-3.
//
// Stands in for Sai2Model::Sai2Model(urdf) (un-vendored sai2-model over RBDL + sai2-urdfreader/tinyxml2)
// as the reference uses it at examples/04-task_and_redundancy/04-task_and_redundancy.cpp:67 and for the world file at
// examples/04-task_and_redundancy/world.urdf:3,5-7 (gravity attribute of <world>, optional <origin> of the <robot> element =
// T_world_robot): parse the XML, weld the root link to the world, merge links attached through fixed joints into their movable
// parent (mass, centre of mass, inertia combined; the link stays addressable by name as a frame fixed in that body), take the
// initial posture from <calibration falling="deg"> (revolute) / rising (prismatic), the limits from <limit>.
// Only serial chains are accepted.  Host code only; compiled into libosc_b200.so next to the C ABI.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/osc_b200.h"

namespace osc {
void set_error(const std::string& msg);
}

namespace {

// ---------------------------------------------------------------------------------------------------------------------
// a minimal XML reader: elements, attributes, comments, declarations; text content is ignored
// ---------------------------------------------------------------------------------------------------------------------
struct XmlNode {
	std::string name;
	std::map<std::string, std::string> attr;
	std::vector<std::unique_ptr<XmlNode>> kids;
	const XmlNode* child(const char* n) const {
		for (auto& k : kids)
			if (k->name == n) return k.get();
		return nullptr;
	}
	const char* get(const char* a) const {
		auto it = attr.find(a);
		return it == attr.end() ? nullptr : it->second.c_str();
	}
};

struct XmlParser {
	const std::string& s;
	size_t i = 0;
	std::string err;
	explicit XmlParser(const std::string& text) : s(text) {}
	void skip_ws() { while (i < s.size() && isspace((unsigned char)s[i])) ++i; }
	bool starts(const char* t) const { return s.compare(i, strlen(t), t) == 0; }
	bool skip_misc() {  // whitespace, text, comments, <? ?>, <! >
		for (;;) {
			while (i < s.size() && s[i] != '<') ++i;
			if (i >= s.size()) return false;
			if (starts("<!--")) {
				size_t e = s.find("-->", i + 4);
				if (e == std::string::npos) { err = "unterminated comment"; return false; }
				i = e + 3;
			} else if (starts("<?")) {
				size_t e = s.find("?>", i + 2);
				if (e == std::string::npos) { err = "unterminated declaration"; return false; }
				i = e + 2;
			} else if (starts("<!")) {
				size_t e = s.find('>', i + 2);
				if (e == std::string::npos) { err = "unterminated <!"; return false; }
				i = e + 1;
			} else {
				return true;
			}
		}
	}
	std::string name_token() {
		size_t b = i;
		while (i < s.size() && (isalnum((unsigned char)s[i]) || s[i] == '_' || s[i] == '-' || s[i] == ':' || s[i] == '.')) ++i;
		return s.substr(b, i - b);
	}
	std::unique_ptr<XmlNode> element() {  // at '<' of an opening tag
		++i;
		std::unique_ptr<XmlNode> n(new XmlNode);
		n->name = name_token();
		if (n->name.empty()) { err = "malformed tag"; return nullptr; }
		for (;;) {
			skip_ws();
			if (i >= s.size()) { err = "unterminated tag <" + n->name; return nullptr; }
			if (starts("/>")) { i += 2; return n; }
			if (s[i] == '>') { ++i; break; }
			std::string a = name_token();
			skip_ws();
			if (a.empty() || i >= s.size() || s[i] != '=') { err = "malformed attribute in <" + n->name; return nullptr; }
			++i;
			skip_ws();
			if (i >= s.size() || (s[i] != '"' && s[i] != '\'')) { err = "unquoted attribute in <" + n->name; return nullptr; }
			const char q = s[i++];
			size_t e = s.find(q, i);
			if (e == std::string::npos) { err = "unterminated attribute value in <" + n->name; return nullptr; }
			n->attr[a] = s.substr(i, e - i);
			i = e + 1;
		}
		for (;;) {  // children until the closing tag
			if (!skip_misc()) { if (err.empty()) err = "missing </" + n->name + ">"; return nullptr; }
			if (starts("</")) {
				i += 2;
				std::string c = name_token();
				skip_ws();
				if (c != n->name || i >= s.size() || s[i] != '>') { err = "mismatched </" + c + "> for <" + n->name + ">"; return nullptr; }
				++i;
				return n;
			}
			std::unique_ptr<XmlNode> k = element();
			if (!k) return nullptr;
			n->kids.push_back(std::move(k));
		}
	}
	std::unique_ptr<XmlNode> document() {
		if (!skip_misc()) { if (err.empty()) err = "no root element"; return nullptr; }
		return element();
	}
};

bool read_file(const char* path, std::string& out) {
	FILE* f = fopen(path, "rb");
	if (!f) return false;
	char buf[65536];
	size_t n;
	while ((n = fread(buf, 1, sizeof buf, f)) > 0) out.append(buf, n);
	fclose(f);
	return true;
}

// ---------------------------------------------------------------------------------------------------------------------
// small 3-vector / 3x3 helpers (row-major)
// ---------------------------------------------------------------------------------------------------------------------
struct V { double v[3] = {0, 0, 0}; };
struct M { double m[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}; };
M mul(const M& a, const M& b) {
	M r;
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) r.m[i][j] = a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j] + a.m[i][2] * b.m[2][j];
	return r;
}
M tr(const M& a) {
	M r;
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) r.m[i][j] = a.m[j][i];
	return r;
}
V mul(const M& a, const V& x) {
	V r;
	for (int i = 0; i < 3; ++i) r.v[i] = a.m[i][0] * x.v[0] + a.m[i][1] * x.v[1] + a.m[i][2] * x.v[2];
	return r;
}
V add(const V& a, const V& b) { V r; for (int i = 0; i < 3; ++i) r.v[i] = a.v[i] + b.v[i]; return r; }
// URDF fixed-axis roll / pitch / yaw: R = Rz(yaw) Ry(pitch) Rx(roll)
M rpy(const V& a) {
	const double cr = cos(a.v[0]), sr = sin(a.v[0]), cp = cos(a.v[1]), sp = sin(a.v[1]), cy = cos(a.v[2]), sy = sin(a.v[2]);
	M Rx, Ry, Rz;
	Rx.m[1][1] = cr; Rx.m[1][2] = -sr; Rx.m[2][1] = sr; Rx.m[2][2] = cr;
	Ry.m[0][0] = cp; Ry.m[0][2] = sp; Ry.m[2][0] = -sp; Ry.m[2][2] = cp;
	Rz.m[0][0] = cy; Rz.m[0][1] = -sy; Rz.m[1][0] = sy; Rz.m[1][1] = cy;
	return mul(mul(Rz, Ry), Rx);
}
// up to 3 numbers of an attribute; a malformed token (the sliding-base URDF has "0.-75") reads as 0
V floats3(const char* s) {
	V r;
	if (!s) return r;
	int k = 0;
	const char* p = s;
	while (*p && k < 3) {
		while (*p && isspace((unsigned char)*p)) ++p;
		if (!*p) break;
		const char* e = p;
		while (*e && !isspace((unsigned char)*e)) ++e;
		std::string tok(p, e - p);
		char* endp = nullptr;
		const double v = strtod(tok.c_str(), &endp);
		r.v[k++] = (endp && *endp == 0) ? v : 0.0;
		p = e;
	}
	return r;
}
double num(const char* s, double dflt = 0.0) { return s ? atof(s) : dflt; }
void origin_of(const XmlNode* e, V& p, M& R) {
	p = V(); R = M();
	const XmlNode* o = e ? e->child("origin") : nullptr;
	if (!o) return;
	p = floats3(o->get("xyz"));
	R = rpy(floats3(o->get("rpy")));
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// URDF -> chain
// ---------------------------------------------------------------------------------------------------------------------
struct UrdfLink { double mass = 0; V com; M I; bool has_inertia = false; };
struct UrdfJoint {
	std::string name, type, parent, child;
	V xyz, axis; M rot;
	double lower = 0, upper = 0, effort = 0, velocity = 0, q0 = 0;
};
struct LinkFrame { std::string name; int body; V xyz; M rot; };
struct Part { double m; V c; M I; };

struct osc_urdf {
	osc_chain_desc chain;
	double q0[OSC_MAX_DOF];
	std::vector<LinkFrame> links;
	std::vector<std::string> joint_names;
	std::string robot_name;
};

namespace {

struct Builder {
	std::map<std::string, UrdfLink> links;
	std::vector<std::string> link_order;
	std::vector<UrdfJoint> joints;
	std::vector<UrdfJoint> chain_joints;
	std::vector<std::vector<Part>> parts;
	std::vector<LinkFrame> frames;
	std::string err;

	bool visit(const std::string& link, int body, const V& p_in_body, const M& R_in_body) {
		frames.push_back(LinkFrame{link, body, p_in_body, R_in_body});
		if (body >= 0) {
			const UrdfLink& L = links[link];
			if (L.mass > 0) parts[body].push_back(Part{L.mass, add(p_in_body, mul(R_in_body, L.com)), mul(mul(R_in_body, L.I), tr(R_in_body))});
		}
		for (const UrdfJoint& j : joints) {
			if (j.parent != link) continue;
			const V p = add(p_in_body, mul(R_in_body, j.xyz));
			const M R = mul(R_in_body, j.rot);
			if (j.type == "fixed") {
				if (!visit(j.child, body, p, R)) return false;
			} else if (j.type == "revolute" || j.type == "continuous" || j.type == "prismatic") {
				if ((int)chain_joints.size() != body + 1) { err = "not a serial chain (branching at link " + link + ")"; return false; }
				if ((int)chain_joints.size() >= OSC_MAX_DOF) { err = "chain has more than OSC_MAX_DOF joints"; return false; }
				UrdfJoint c = j;
				const double na = sqrt(j.axis.v[0] * j.axis.v[0] + j.axis.v[1] * j.axis.v[1] + j.axis.v[2] * j.axis.v[2]);
				if (!(na > 0)) { err = "joint " + j.name + " has a zero axis"; return false; }
				for (int a = 0; a < 3; ++a) c.axis.v[a] = j.axis.v[a] / na;
				c.xyz = p; c.rot = R;
				chain_joints.push_back(c);
				parts.emplace_back();
				if (!visit(j.child, (int)chain_joints.size() - 1, V(), M())) return false;
			} else {
				err = "unsupported joint type " + j.type;
				return false;
			}
		}
		return true;
	}
};

int fail(const std::string& msg) {
	osc::set_error(msg);
	return OSC_ERR_INVALID_ARGUMENT;
}

}  // namespace

extern "C" {
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

int osc_urdf_load(const char* urdf_path, const char* world_path, const char* robot_name, osc_urdf_t** out) {
	if (!urdf_path || !out) return fail("null argument in osc_urdf_load");
	*out = nullptr;
	std::string text;
	if (!read_file(urdf_path, text)) return fail(std::string("cannot read URDF file ") + urdf_path);
	XmlParser xp(text);
	std::unique_ptr<XmlNode> root = xp.document();
	if (!root) return fail(std::string("URDF parse error in ") + urdf_path + ": " + xp.err);
	if (root->name != "robot") return fail("URDF root element must be <robot>");
	Builder b;
	std::map<std::string, bool> is_child;
	for (auto& e : root->kids) {
		if (e->name == "link") {
			const char* nm = e->get("name");
			if (!nm) return fail("URDF <link> without a name");
			UrdfLink L;
			if (const XmlNode* in = e->child("inertial")) {
				M Rc;
				origin_of(in, L.com, Rc);
				const XmlNode* me = in->child("mass");
				const XmlNode* ie = in->child("inertia");
				L.mass = me ? num(me->get("value")) : 0.0;
				if (ie) {
					M I;
					I.m[0][0] = num(ie->get("ixx")); I.m[0][1] = I.m[1][0] = num(ie->get("ixy")); I.m[0][2] = I.m[2][0] = num(ie->get("ixz"));
					I.m[1][1] = num(ie->get("iyy")); I.m[1][2] = I.m[2][1] = num(ie->get("iyz")); I.m[2][2] = num(ie->get("izz"));
					L.I = mul(mul(Rc, I), tr(Rc));
				} else {
					L.I = M(); for (int i = 0; i < 3; ++i) L.I.m[i][i] = 0;
				}
				L.has_inertia = true;
			} else {
				for (int i = 0; i < 3; ++i) L.I.m[i][i] = 0;
			}
			b.links[nm] = L;
			b.link_order.push_back(nm);
		} else if (e->name == "joint") {
			UrdfJoint j;
			const char* nm = e->get("name"); const char* ty = e->get("type");
			const XmlNode* pa = e->child("parent"); const XmlNode* ch = e->child("child");
			if (!nm || !ty || !pa || !ch || !pa->get("link") || !ch->get("link")) return fail("URDF <joint> needs name, type, <parent link>, <child link>");
			j.name = nm; j.type = ty; j.parent = pa->get("link"); j.child = ch->get("link");
			origin_of(e.get(), j.xyz, j.rot);
			const XmlNode* ax = e->child("axis");
			if (ax) j.axis = floats3(ax->get("xyz")); else { j.axis = V(); j.axis.v[0] = 1.0; }
			if (const XmlNode* lim = e->child("limit")) {
				j.lower = num(lim->get("lower")); j.upper = num(lim->get("upper")); j.effort = num(lim->get("effort")); j.velocity = num(lim->get("velocity"));
			}
			if (const XmlNode* cal = e->child("calibration")) {  // initial posture: degrees for revolute joints, metres for prismatic ones
				if (j.type == "prismatic") j.q0 = num(cal->get("rising") ? cal->get("rising") : cal->get("falling"));
				else j.q0 = num(cal->get("falling") ? cal->get("falling") : cal->get("rising")) * (M_PI / 180.0);
			}
			is_child[j.child] = true;
			b.joints.push_back(j);
		}
	}
	std::string root_link;
	int nroots = 0;
	for (const std::string& nm : b.link_order)
		if (!is_child.count(nm)) { root_link = nm; ++nroots; }
	if (nroots != 1) return fail("URDF must have exactly one root link");
	for (const UrdfJoint& j : b.joints)
		if (!b.links.count(j.parent) || !b.links.count(j.child)) return fail("joint " + j.name + " refers to an unknown link");
	if (!b.visit(root_link, -1, V(), M())) return fail(b.err);
	const int n = (int)b.chain_joints.size();
	if (n < 1) return fail("URDF has no movable joint");

	std::unique_ptr<osc_urdf> u(new osc_urdf);
	memset(&u->chain, 0, sizeof u->chain);
	memset(u->q0, 0, sizeof u->q0);
	u->robot_name = root->get("name") ? root->get("name") : "robot";
	osc_chain_desc& c = u->chain;
	c.n = n;
	for (int i = 0; i < n; ++i) {
		const UrdfJoint& j = b.chain_joints[i];
		c.joint_type[i] = j.type == "prismatic" ? OSC_JOINT_PRISMATIC : OSC_JOINT_REVOLUTE;
		for (int a = 0; a < 3; ++a) { c.joint_axis[i][a] = j.axis.v[a]; c.joint_xyz[i][a] = j.xyz.v[a]; }
		for (int k = 0; k < 9; ++k) c.joint_rot[i][k] = j.rot.m[k / 3][k % 3];
		c.q_lower[i] = j.lower; c.q_upper[i] = j.upper; c.effort[i] = j.effort;
		u->q0[i] = j.q0;
		u->joint_names.push_back(j.name);
		// merged body: total mass, centre of mass, inertia about it (parallel-axis theorem per part)
		double m = 0; V cm;
		for (const Part& p : b.parts[i]) { m += p.m; for (int a = 0; a < 3; ++a) cm.v[a] += p.m * p.c.v[a]; }
		if (m > 0) {
			for (int a = 0; a < 3; ++a) cm.v[a] /= m;
			double I[3][3] = {{0}};
			for (const Part& p : b.parts[i]) {
				const double d[3] = {p.c.v[0] - cm.v[0], p.c.v[1] - cm.v[1], p.c.v[2] - cm.v[2]};
				const double dd = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
				for (int r = 0; r < 3; ++r)
					for (int s = 0; s < 3; ++s) I[r][s] += p.I.m[r][s] + p.m * ((r == s ? dd : 0.0) - d[r] * d[s]);
			}
			c.mass[i] = m;
			for (int a = 0; a < 3; ++a) c.com[i][a] = cm.v[a];
			for (int k = 0; k < 9; ++k) c.inertia[i][k] = I[k / 3][k % 3];
		}
	}
	// world: identity pose, gravity (0, 0, -9.81) unless the world file says otherwise
	const double Id[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
	memcpy(c.world_rot, Id, sizeof Id);
	c.gravity[2] = -9.81;
	if (world_path) {
		std::string wt;
		if (!read_file(world_path, wt)) return fail(std::string("cannot read world file ") + world_path);
		XmlParser wp(wt);
		std::unique_ptr<XmlNode> w = wp.document();
		if (!w) return fail(std::string("world file parse error in ") + world_path + ": " + wp.err);
		if (w->name != "world") return fail("world file root element must be <world>");
		if (w->get("gravity")) { const V g = floats3(w->get("gravity")); for (int a = 0; a < 3; ++a) c.gravity[a] = g.v[a]; }
		const XmlNode* rb = nullptr;
		for (auto& e : w->kids)
			if (e->name == "robot" && (!robot_name || (e->get("name") && !strcmp(e->get("name"), robot_name)))) { rb = e.get(); break; }
		if (!rb) return fail(robot_name ? std::string("world file has no <robot name=\"") + robot_name + "\">" : std::string("world file has no <robot>"));
		V p; M R;
		origin_of(rb, p, R);  // T_world_robot
		for (int a = 0; a < 3; ++a) c.world_xyz[a] = p.v[a];
		for (int k = 0; k < 9; ++k) c.world_rot[k] = R.m[k / 3][k % 3];
	}
	u->links = b.frames;
	*out = u.release();
	return OSC_OK;
}

void osc_urdf_free(osc_urdf_t* u) { delete u; }

int osc_urdf_chain(const osc_urdf_t* u, osc_chain_desc* chain, double* q0) {
	if (!u || !chain) return fail("null argument in osc_urdf_chain");
	*chain = u->chain;
	if (q0) memcpy(q0, u->q0, sizeof(double) * u->chain.n);
	return OSC_OK;
}
int osc_urdf_link_count(const osc_urdf_t* u) { return u ? (int)u->links.size() : 0; }
const char* osc_urdf_link_name(const osc_urdf_t* u, int i) { return (u && i >= 0 && i < (int)u->links.size()) ? u->links[i].name.c_str() : nullptr; }
const char* osc_urdf_joint_name(const osc_urdf_t* u, int i) { return (u && i >= 0 && i < (int)u->joint_names.size()) ? u->joint_names[i].c_str() : nullptr; }
int osc_urdf_link(const osc_urdf_t* u, const char* link_name, int* body, double xyz[3], double rot[9]) {
	if (!u || !link_name) return fail("null argument in osc_urdf_link");
	for (const LinkFrame& f : u->links)
		if (f.name == link_name) {
			if (body) *body = f.body;
			if (xyz) for (int a = 0; a < 3; ++a) xyz[a] = f.xyz.v[a];
			if (rot) for (int k = 0; k < 9; ++k) rot[k] = f.rot.m[k / 3][k % 3];
			return OSC_OK;
		}
	return fail(std::string("no link named ") + link_name + " in the URDF");
}

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
}  // extern "C"
