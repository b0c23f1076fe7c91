#include "lattice_hop.h"
int hop(float t, float tt, float dt, float* out) { return ngp_lattice_hop(t, tt, dt, out); }
float loop(float t, float tt, float dt) { do { t += dt; } while (t < tt); return t; }
/* returns number of mismatches; counts fast-path uses in *used */
long run(long n, unsigned long long seed, float dt, float tmax, long* used) {
	long bad = 0; *used = 0;
	for (long i = 0; i < n; ++i) {
		seed = seed * 6364136223846793005ULL + 1442695040888963407ULL; unsigned a = seed >> 33;
		seed = seed * 6364136223846793005ULL + 1442695040888963407ULL; unsigned b = seed >> 33;
		seed = seed * 6364136223846793005ULL + 1442695040888963407ULL; unsigned c = seed >> 33;
		float t = 0.05f + (a / 2147483648.0f) * tmax;
		float dist = (b / 2147483648.0f) * ((c & 7) == 0 ? 0.3f : 0.012f);
		if ((c & 63) == 1) dist = 0.f;
		/* put targets exactly on / next to lattice points now and then */
		float tt = t + dist;
		if ((c & 15) == 2) { float x = t; int m = (c >> 8) % 9 + 1; for (int q = 0; q < m; ++q) x += dt; tt = x; if (c & 0x10000) tt = nextafterf(x, 0.f); else if (c & 0x20000) tt = nextafterf(x, 1e9f); }
		float w = loop(t, tt, dt), g;
		if (hop(t, tt, dt, &g)) { ++*used; if (memcmp(&w, &g, 4)) ++bad; }
	}
	return bad;
}

/* the marching loop's use of the cached form: many hops and single steps in a row from one start, constants re-entered exactly
   where the kernel re-enters them; returns mismatches against the plain loop, *used counts closed-form hops */
long run_cached(long n_rays, int hops_per_ray, unsigned long long seed, float dt, float tmax, long* used) {
	long bad = 0; *used = 0;
	for (long i = 0; i < n_rays; ++i) {
		seed = seed * 6364136223846793005ULL + 1442695040888963407ULL; unsigned a = seed >> 33;
		float t = 0.05f + (a / 2147483648.0f) * tmax, ref = t;
		ngp_lattice L; ngp_lattice_enter(&L, t, dt, 1);
		for (int h = 0; h < hops_per_ray; ++h) {
			seed = seed * 6364136223846793005ULL + 1442695040888963407ULL; unsigned b = seed >> 33;
			seed = seed * 6364136223846793005ULL + 1442695040888963407ULL; unsigned c = seed >> 33;
			if ((c & 3) == 0) {                       /* a sample: one plain step */
				t += dt; ref += dt;
				if (!(t < L.top)) ngp_lattice_enter(&L, t, dt, 1);
			} else {
				float dist = (b / 2147483648.0f) * ((c & 28) == 0 ? 0.25f : 0.012f);
				float tt = t + dist;
				if ((c & 96) == 32) { float x = t; int m = (c >> 8) % 9 + 1; for (int q = 0; q < m; ++q) x += dt; tt = (c & 0x10000) ? nextafterf(x, 0.f) : (c & 0x20000) ? nextafterf(x, 1e9f) : x; }
				ref = loop(ref, tt, dt);
				float g;
				if (ngp_lattice_hop_cached(&L, t, tt, &g)) { t = g; ++*used; }
				else { t = loop(t, tt, dt); ngp_lattice_enter(&L, t, dt, 1); }
			}
			if (memcmp(&t, &ref, 4)) { ++bad; t = ref; ngp_lattice_enter(&L, t, dt, 1); }
		}
	}
	return bad;
}
