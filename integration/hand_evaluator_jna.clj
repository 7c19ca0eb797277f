;; Drop-in for namespace montecarlo.hand-evaluator over libmcpoker.so (include/mcpoker.h) through JNA.
;; NOT EXECUTABLE ON THIS BOX (no JVM in the image); shipped as the binding a maintainer of the reference
;; would add.  See INTEGRATION.md for the ABI map.  Same public vars as the reference namespace; the
;; value-producing ones go to the GPU library, the small predicates (test surface only) are host code.
(ns montecarlo.hand-evaluator
  (:require [montecarlo.database :as database])
  (:import [com.sun.jna Function NativeLibrary Memory]
           [com.sun.jna.ptr IntByReference]))

;; ---- native plumbing ---------------------------------------------------------------------------------
(def ^:private lib (delay (NativeLibrary/getInstance "mcpoker")))
(defn- f ^Function [nm] (.getFunction ^NativeLibrary @lib nm))
(defn- call-int [nm & args] (.invokeInt (f nm) (object-array args)))
(defn- check [rc]
  (when (neg? rc)
    (let [msg (.invokeString (f "mc_last_error") (object-array 0) false)]
      (if (= rc -5)
        (throw (clojure.lang.ArityException. 0 "montecarlo.hand-evaluator/max-hand-value"))
        (throw (ex-info msg {:code rc})))))
  rc)
(defonce ^:private started (delay (check (call-int "mc_init" (int 0) (int 0)))))
(defn shutdown! [] (.invokeVoid (f "mc_shutdown") (object-array 0)))

;; ---- packing -----------------------------------------------------------------------------------------
(defn- rank-of [c] (long (:rank c)))
(defn- suit-ids [cards]
  (let [ids (zipmap (distinct (map :suit cards)) (range))]
    (when (> (count ids) 4) (throw (ex-info "more than four distinct suits" {})))
    (mapv (comp ids :suit) cards)))

(defn- resuit-repeats
  "A card set cannot hold the same card twice; when the hand is not single-suited a repeated card may move to
   any free suit without changing its value (only all-suits-equal is ever asked)."
  [sids ranks]
  (if (= (count (distinct (map vector sids ranks))) (count sids))
    sids
    (do (when (< (count (distinct sids)) 2) (throw (ex-info "repeated card in a single-suited hand" {})))
        (loop [i 0 seen #{} out sids]
          (if (= i (count out))
            out
            (let [card [(out i) (ranks i)]]
              (if (seen card)
                (let [later (set (map vector (subvec out (inc i)) (subvec ranks (inc i))))
                      free  (first (remove #(or (seen [% (ranks i)]) (later [% (ranks i)])) (range 4)))]
                  (when (nil? free) (throw (ex-info "more than four cards of one rank" {})))
                  (recur (inc i) (conj seen [free (ranks i)]) (assoc out i free)))
                (recur (inc i) (conj seen card) out))))))))

(defn- eval-key [cards]
  @started
  (let [cards (vec cards) n (count cards)
        ranks (mapv rank-of cards)
        sids  (resuit-repeats (suit-ids cards) ranks)
        s (Memory. (* 4 (max n 1))) r (Memory. (* 4 (max n 1))) k (IntByReference.)]
    (dotimes [i n] (.setInt s (* 4 i) (int (sids i))) (.setInt r (* 4 i) (int (ranks i))))
    (check (call-int "mc_eval_cards" s r (int n) k))
    (Integer/toUnsignedLong (.getValue k))))

(def ^:private hit-len [5 2 4 3 5 5 5 4 5])
(defn key->value [k]
  (let [cat (bit-shift-right k 20)
        rs  (mapv #(bit-and 15 (bit-shift-right k (- 16 (* 4 %)))) (range 5))
        h   (hit-len cat)]
    [cat (subvec rs 0 h) (subvec rs h)]))

;; ---- plain predicates (test surface) -------------------------------------------------------------------
(defn constant-coll? [coll] (when (apply = (or (seq coll) [nil])) coll))
(defn constant-cards? [cards att] (when (constant-coll? (map att cards)) cards))
(defn high-card [a b] (if (pos? (compare (:rank a) (:rank b))) a b))
(defn high-hand [hand-a hand-b]
  (let [top-a (reduce high-card hand-a) top-b (reduce high-card hand-b)]
    (if (= top-a (high-card top-a top-b)) hand-a hand-b)))
(defn straight? [cards]
  (let [rs (sort (map :rank cards))]
    (when (every? (fn [[x y]] (= (inc x) y)) (partition 2 1 rs)) cards)))
(defn flush? [cards] (constant-cards? cards :suit))
(defn- subsets [xs k]
  (cond (zero? k) [[]] (empty? xs) []
        :else (concat (map #(cons (first xs) %) (subsets (rest xs) (dec k))) (subsets (rest xs) k))))
(defn n-of-a-kinds [cards n] (filter #(constant-cards? % :rank) (subsets (seq cards) n)))
(defn n-of-a-kind? [cards n] (let [ks (n-of-a-kinds cards n)] (when (seq ks) (reduce high-hand ks))))
(defn four-of-a-kind? [cards] (n-of-a-kind? cards 4))
(defn three-of-a-kind? [cards] (n-of-a-kind? cards 3))
(defn pair? [cards] (n-of-a-kind? cards 2))
;; n-pairs? :71-79.  The reference walks the 120 (or 24, ...) index-lexicographic permutations of the cards and keeps
;; the first whose leading k1 cards and following k2 cards are each of one rank.  `take` past the end of the
;; permutation gives a SHORT (or empty) group, and a group of fewer than two cards is rank-constant, so four 2s are
;; a "full house" (test :36: trips + the one card left) and a "two pair" (:49).  Same search here, so the same first
;; hit and the same truthiness for every input; only the values of `evaluator` go to the GPU.
(defn- orderings [v]                      ; index-lexicographic permutations of a vector
  (if (empty? v)
    [()]
    (for [i (range (count v))
          tail (orderings (into (subvec v 0 i) (subvec v (inc i))))]
      (cons (v i) tail))))
(defn n-pairs? [cards k1 k2]
  (let [one-rank? #(constant-cards? % :rank)]
    (first (for [o (orderings (vec cards))
                 :let [g1 (sort-by :rank (take k1 o))
                       g2 (sort-by :rank (take k2 (drop k1 o)))]
                 :when (and (one-rank? g1) (one-rank? g2))]
             [g1 g2]))))
(defn full-house? [cards] (n-pairs? cards 3 2))
(defn two-pair? [cards] (n-pairs? cards 2 2))
(defn cards->ranks [cards] (reverse (sort (map :rank cards))))

;; ---- GPU-backed values ---------------------------------------------------------------------------------
(defn evaluator [cards] (key->value (eval-key cards)))
(defn max-hand-value [val-a val-b] (if (pos? (compare val-a val-b)) val-a val-b))
(defn player->hand-value [board player-id]
  (let [pocket    @(database/player-hand player-id (:room board))
        available (concat @(:community-cards board) pocket)]
    (key->value (eval-key available))))

;; ---- additive: many hands in one call -------------------------------------------------------------------
;; mc_eval_batch_packed48: 6 bytes per hand (seven 6-bit card positions 16*suit + rank, 0 = no card).  The call is
;; bound by the host link, so the compact format is 4/3 of the mask format's throughput.  For large batches take the
;; buffers from mc_alloc_host (page-locked) instead of Memory.
(defn eval-batch
  "hands: a seq of 5..7-card seqs -> vector of [cat hit kick] values, one per hand"
  [hands]
  @started
  (let [hands (vec hands) n (count hands)
        in (Memory. (* 6 (max n 1))) out (Memory. (* 4 (max n 1)))]
    (dotimes [i n]
      (let [cards (vec (hands i))
            ranks (mapv rank-of cards)
            sids  (resuit-repeats (suit-ids cards) ranks)
            h     (reduce (fn [acc k] (bit-or acc (bit-shift-left (+ (* 16 (long (sids k))) (long (ranks k))) (* 6 k))))
                          0 (range (count cards)))]
        (.setInt in (* 6 i) (unchecked-int h))
        (.setShort in (+ 4 (* 6 i)) (unchecked-short (bit-shift-right h 32)))))
    (check (call-int "mc_eval_batch_packed48" in (long n) out))
    (mapv #(key->value (Integer/toUnsignedLong (.getInt out (* 4 %)))) (range n))))
