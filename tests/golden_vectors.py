"""Golden vectors held by the reference's own C-API tests (compared there with relative 1e-3,
cpp/tests/c_api/test_utils.cpp:21-27)."""
import numpy as np

# cpp/tests/c_api/pagerank_test.c:211-249 (and katz_test.c:103-129 on the same graph)
G6_SRC = np.array([0, 1, 1, 2, 2, 2, 3, 4], np.int32)
G6_DST = np.array([1, 3, 4, 0, 1, 3, 5, 5], np.int32)
G6_WGT = np.array([0.1, 2.1, 1.1, 5.1, 3.1, 4.1, 7.2, 3.2], np.float32)
G6_PAGERANK = dict(alpha=0.95, epsilon=1e-4, max_iterations=20,
                   result=np.array([0.0915528, 0.168382, 0.0656831, 0.191468, 0.120677, 0.362237]))
G6_KATZ = dict(alpha=0.01, beta=1.0, epsilon=1e-6, max_iterations=1000,
               result=np.array([0.410614, 0.403211, 0.390689, 0.415175, 0.395125, 0.433226]))

# cpp/tests/c_api/pagerank_test.c:251-287; also the doctest of pylibcugraph/pagerank.pyx:135-146
P4_SRC = np.array([0, 1, 2], np.int32)
P4_DST = np.array([1, 2, 3], np.int32)
P4_WGT = np.array([1.0, 1.0, 1.0], np.float32)
P4_PAGERANK = dict(alpha=0.85, epsilon=1e-6, max_iterations=500,
                   result=np.array([0.11615584790706635, 0.21488840878009796,
                                    0.29881080985069275, 0.37014490365982056]))
# cpp/tests/c_api/pagerank_test.c:289-318
P4_PERSONALIZED = dict(alpha=0.85, epsilon=1e-6, max_iterations=500,
                       vertices=np.array([0, 1, 2, 3], np.int32),
                       values=np.array([0.1, 0.2, 0.3, 0.4], np.float32),
                       result=np.array([0.0559233, 0.159381, 0.303244, 0.481451]))

# cpp/tests/c_api/eigenvector_centrality_test.c:93-110
E6_SRC = np.array([0, 1, 1, 2, 2, 2, 3, 4, 1, 3, 4, 0, 1, 3, 5, 5], np.int32)
E6_DST = np.array([1, 3, 4, 0, 1, 3, 5, 5, 0, 1, 1, 2, 2, 2, 3, 4], np.int32)
E6_WGT = np.array([0.1, 2.1, 1.1, 5.1, 3.1, 4.1, 7.2, 3.2] * 2, np.float32)
E6_EIGENVECTOR = dict(epsilon=1e-6, max_iterations=200,
                      result=np.array([0.236325, 0.292055, 0.458457, 0.60533, 0.190498, 0.495942]))


def nearly_equal(a, b, eps=1e-3):
    """cpp/tests/c_api/test_utils.cpp:21-27"""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.all(np.abs(a - b) <= eps * np.maximum(np.abs(a), np.abs(b)))


# cpp/tests/c_api/hits_test.c:155-185 (test_hits) and :221-253 (test_hits_with_initial): graph G6,
# normalize = FALSE (scores are max-normalised), epsilon 1e-4, 20 iterations at most
G6_HITS = dict(epsilon=0.0001, max_iterations=20,
               hubs=np.array([0.347296, 0.532089, 1, 0.00003608, 0.00003608, 0]),
               authorities=np.array([0.652703, 0.879385, 0, 1, 0.347296, 0.00009136]))
G6_HITS_INITIAL = dict(epsilon=0.0001, max_iterations=20,
                       initial_vertices=np.array([0, 1, 2, 3, 4], dtype=np.int32),
                       initial_hubs=np.array([0.347296, 0.532089, 1, 0.00003608, 0.00003608]),
                       hubs=np.array([0.347296, 0.532089, 1, 0.00000959, 0.00000959, 0]),
                       authorities=np.array([0.652704, 0.879385, 0, 1, 0.347296, 0.00002428]))
# cpp/tests/c_api/hits_test.c:255-300 (test_hits_bigger): 54 vertices, 48 edges, epsilon 1e-6,
# 100 iterations at most, normalize = FALSE
H54_SRC = np.array([29, 45, 6, 8, 16, 45, 8, 16, 6, 38, 45, 45, 48, 45, 45, 45,
                    45, 48, 53, 45, 6, 45, 38, 45, 38, 45, 16, 45, 38, 16, 45, 45,
                    38, 6, 38, 45, 45, 45, 16, 38, 6, 45, 29, 45, 29, 6, 38, 6], dtype=np.int32)
H54_DST = np.array([45, 45, 16, 45, 6, 45, 45, 16, 45, 38, 45, 6, 45, 38, 16, 45,
                    45, 45, 45, 53, 29, 16, 45, 8, 8, 16, 45, 38, 45, 6, 45, 45,
                    6, 6, 16, 38, 16, 45, 45, 6, 16, 6, 53, 16, 38, 45, 45, 16], dtype=np.int32)
_h54_hubs = np.zeros(54)
_h54_hubs[[6, 8, 16, 29, 38, 45, 48, 53]] = [0.323569, 0.156401, 0.253312, 0.110617, 0.365733, 1,
                                             0.156401, 0.0782005]
_h54_auth = np.zeros(54)
_h54_auth[[6, 8, 16, 29, 38, 45, 53]] = [0.321874, 0.123424, 0.595522, 0.0292397, 0.314164, 1,
                                         0.100368]
H54_HITS = dict(epsilon=0.000001, max_iterations=100, hubs=_h54_hubs, authorities=_h54_auth)

# cpp/tests/centrality/legacy/katz_centrality_test.cu:114-183 with datasets/ref/katz/*.csv: the ten
# best vertices (best first) of Katz centrality on the UNWEIGHTED graph, alpha = 1 / (max out-degree
# + 1), 100 iterations at most, tolerance 1e-6, normalised.  (netscience is commented out there.)
KATZ_TOP10 = {
    "karate.mtx": [33, 0, 32, 2, 1, 31, 3, 8, 13, 23],
    "polbooks.mtx": [8, 12, 84, 3, 73, 72, 66, 30, 11, 47],
    "dolphins.mtx": [14, 37, 45, 33, 51, 29, 20, 40, 50, 38],
}

# python/pylibcugraph/pylibcugraph/tests/test_pagerank.py:22-77: alpha 0.85, epsilon 1e-6, 500
# iterations at most, float32, graphs made with renumber=False; compared with
# pytest.approx(expected, 1e-4) (relative).  karate.csv / dolphins.csv hold the same directed edge
# sets as the symmetric karate.mtx / dolphins.mtx expanded (unit weights).  "Simple_1" is P4 above;
# "Simple_2" is the G6 graph.
PLC_PAGERANK = dict(alpha=0.85, epsilon=1.0e-6, max_iterations=500, rel=1e-4)
PLC_PAGERANK_RESULTS = {
    "karate.mtx": np.array(
        [0.096998, 0.052877, 0.057078, 0.035860, 0.021978, 0.029111, 0.029111, 0.024491, 0.029766,
         0.014309, 0.021978, 0.009565, 0.014645, 0.029536, 0.014536, 0.014536, 0.016784, 0.014559,
         0.014536, 0.019605, 0.014536, 0.014559, 0.014536, 0.031522, 0.021076, 0.021006, 0.015044,
         0.025640, 0.019573, 0.026288, 0.024590, 0.037158, 0.071693, 0.100919]),
    "dolphins.mtx": np.array(
        [0.01696534, 0.02465084, 0.01333804, 0.00962903, 0.00507979, 0.01442816, 0.02005379,
         0.01564308, 0.01709825, 0.02345867, 0.01510835, 0.00507979, 0.0048353, 0.02615709,
         0.03214436, 0.01988301, 0.01662675, 0.03172837, 0.01939547, 0.01292825, 0.02464085,
         0.01693892, 0.00541593, 0.00986347, 0.01690569, 0.01150429, 0.0112102, 0.01713019,
         0.01484573, 0.02645844, 0.0153021, 0.00541593, 0.01330877, 0.02842296, 0.01591988,
         0.00491821, 0.02061337, 0.02987523, 0.02393915, 0.00776477, 0.02196631, 0.01613769,
         0.01761861, 0.02169104, 0.01283079, 0.02951408, 0.00882587, 0.01733948, 0.00526172,
         0.00887672, 0.01923187, 0.03129924, 0.01207255, 0.00818102, 0.02165103, 0.00749415,
         0.0083263, 0.0300956, 0.00496289, 0.01476788, 0.00619018, 0.01103916]),
    "Simple_2": np.array([0.09902544, 0.17307726, 0.0732199, 0.1905103, 0.12379099, 0.34037617]),
}
