"""ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of LoopDetector (PlaceRecognition/LoopDetector.cpp:22-84) and of the
one DBoW3 routine it calls, Vocabulary::score with L1_NORM scoring (DBoW3 is a third-party dependency that is not under
/root/reference and whose version CMakeLists.txt does not pin; restated from DBoW3 src/ScoringObject.cpp, L1Scoring::score,
unchanged since DBoW2).  Pure Python floats = IEEE doubles, accumulated in the order the library walks the two std::maps.
The final std::sort (:79) goes through this image's libstdc++ (oracle/capi.cpp orc_sort_bow_scores).  PARITY UNPINNED
against the libraries themselves (the reference holds no vectors for this path and cannot be built here)."""
import ctypes as C
import sys

import numpy as np


def l1_score(v1, v2):
    """DBoW3 L1Scoring::score(v1, v2); v = (ascending word ids, weights)."""
    w1, x1 = v1
    w2, x2 = v2
    i = j = 0
    s = 0.0
    n1, n2 = len(w1), len(w2)
    while i < n1 and j < n2:
        if w1[i] == w2[j]:
            vi, wi = float(x1[i]), float(x2[j])
            s += abs(vi - wi) - abs(vi) - abs(wi)
            i += 1
            j += 1
        elif w1[i] < w2[j]:
            i += 1  # v1.lower_bound(v2_it->first): same visiting order of the common words
        else:
            j += 1
    return -s / 2.0


class LoopDetector:
    def __init__(self, n_words, interval=10, sort_lib=None):
        self.inverted = [[] for _ in range(n_words)]  # mvInvertedFile: keyframe indices in insertion order
        self.bow, self.ids, self.mark = [], [], []
        self.interval = interval
        self.sort_lib = sort_lib

    def add(self, kf_id, words, weights):  # LoopDetector::add (:22-26)
        k = len(self.bow)
        self.bow.append((np.asarray(words), np.asarray(weights, np.float64)))
        self.ids.append(int(kf_id))
        self.mark.append(-1)  # Frame::mnSharingWord (uninitialised in the reference)
        for w in words:
            self.inverted[int(w)].append(k)
        return k

    def obtain_candidates(self, q, connected):  # :28-84
        connected = set(int(c) for c in connected)
        if not connected:
            return []
        min_score = sys.float_info.max
        for k in sorted(connected):  # the minimum does not depend on the set's (pointer) order
            if k == q:
                continue
            sc = l1_score(self.bow[k], self.bow[q])
            if sc < min_score:
                min_score = sc
        sharing = []
        qid = self.ids[q]
        for w in self.bow[q][0]:
            for k in self.inverted[int(w)]:
                if k == q or self.mark[k] == qid or k in connected:
                    continue
                self.mark[k] = qid
                sharing.append(k)
        loops = []
        for k in sharing:
            sc = l1_score(self.bow[q], self.bow[k])
            if sc > min_score and abs(self.ids[k] - qid) > self.interval:
                loops.append((k, sc))
        if len(loops) > 5:
            idx = np.array([k for k, _ in loops], np.int32)
            sc = np.array([s for _, s in loops], np.float64)
            self.sort_lib.orc_sort_bow_scores(sc.ctypes.data_as(C.c_void_p), idx.ctypes.data_as(C.c_void_p), len(idx))
            loops = [(int(idx[i]), float(sc[i])) for i in range(5)]
        return loops
