// kmeans.mjs — the JS façade of SURVEY.md §8b over the Node-API addon (napi/kmeans_napi.c).
// STATUS: unexecuted — no JS engine exists on the build image.  The Python mirror that IS
// tested (k-means-demo_b200/api.py: init / assign / update / step / run) has the same names,
// argument meaning and error behaviour.
import { createRequire } from 'node:module';
const addon = createRequire(import.meta.url)('./build/Release/kmeans_napi.node');

/** init(points, k, {d = 2, seed} | {d = 2, centroids}) -> state */
export function init(points, k, { d = 2, seed = 0, centroids = null } = {}) {
  const handle = addon.init(points, d, k, centroids, seed);
  return { handle, n: points.length / d, d, k };
}
export const assign = (state) => addon.assign(state.handle);
export const update = (state) => addon.update(state.handle);
export const step = (state) => addon.step(state.handle);

/** run(state, {maxIter}) -> {labels: Int32Array, centroids: Float64Array, iterations, inertia, converged} */
export function run(state, { maxIter = 300 } = {}) {
  const s = addon.run(state.handle, maxIter);
  const labels = new Int32Array(state.n);
  const centroids = new Float64Array(state.k * state.d);
  addon.labels(state.handle, labels);
  addon.centroids(state.handle, centroids);
  return { labels, centroids, iterations: s.iterations, inertia: s.inertia, converged: s.converged };
}
