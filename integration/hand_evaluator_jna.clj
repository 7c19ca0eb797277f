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

;; ---- additive: the compact wire format (4-byte colex ranks in, 2-byte class ids out) ---------------------------
;; mc_eval_batch_colex32: a 7-card hand of the deck travels as the colex rank of its card ids (card.clj:10-13 index),
;; sum_i C(c_i, i+1) with c_0 < ... < c_6; its value comes back as the index in the ascending list of all 7,462
;; five-card values, which mc_class_keys hands out once.  Half the bytes of the 6-byte format on the host link.
(def ^:private suit-index {:hearts 0 :diamonds 1 :spades 2 :clubs 3})
(defn- card-id [c] (+ (* 13 (suit-index (:suit c))) (- (rank-of c) 2)))
(defn- choose [n k] (if (< n k) 0 (reduce (fn [acc i] (quot (* acc (- n i)) (inc i))) 1 (range k))))
(defn colex-rank [cards]
  (reduce + (map-indexed (fn [i id] (choose id (inc i))) (sort (map card-id cards)))))
(def ^:private class-values
  (delay @started
         (let [m (Memory. (* 4 7462))]
           (check (call-int "mc_class_keys" m (int 7462)))
           (mapv #(key->value (Integer/toUnsignedLong (.getInt m (* 4 %)))) (range 7462)))))
(defn eval-batch-ranks
  "hands: a seq of 7-card seqs of deck cards -> vector of [cat hit kick] values"
  [hands]
  @started
  (let [hands (vec hands) n (count hands)
        in (Memory. (* 4 (max n 1))) out (Memory. (* 2 (max n 1)))]
    (dotimes [i n] (.setInt in (* 4 i) (unchecked-int (colex-rank (hands i)))))
    (check (call-int "mc_eval_batch_colex32" in (long n) out))
    (mapv #(@class-values (bit-and 0xFFFF (.getShort out (* 2 %)))) (range n))))

;; ---- additive: end-game in one call (ranking, tie groups, update-stacks; gameplay.clj:122-133) -------------------
(defn- cards->mask [cards]
  (reduce bit-or 0 (map #(bit-shift-left 1 (+ (* 16 (quot (card-id %) 13)) (rem (card-id %) 13) 2)) cards)))
(defn end-game-result
  "community: 3..5 deck cards; pockets: vector of 2-card seqs by seat; pots: seq of {:bet :n :players #{seat ...}}
   -> {:values [...] :groups [group index by seat, 0 = best] :delta [chips won by seat]}"
  [community pockets pots]
  @started
  (let [n (count pockets) k (count pots)
        h (Memory. (* 8 (max n 1))) keys (Memory. 36) grp (Memory. 36) ng (IntByReference.) win (IntByReference.)
        bet (Memory. (* 8 (max k 1))) cnt (Memory. (* 4 (max k 1))) msk (Memory. (* 4 (max k 1))) delta (Memory. 72)]
    (.clear delta)
    (dotimes [i n] (.setLong h (* 8 i) (cards->mask (pockets i))))
    (doseq [[i p] (map-indexed vector pots)]
      (.setLong bet (* 8 i) (long (:bet p))) (.setInt cnt (* 4 i) (int (:n p)))
      (.setInt msk (* 4 i) (int (reduce bit-or 0 (map #(bit-shift-left 1 %) (:players p))))))
    (check (call-int "mc_end_game" (cards->mask community) h (int n) bet cnt msk (int k) keys grp ng win delta))
    {:values (mapv #(key->value (Integer/toUnsignedLong (.getInt keys (* 4 %)))) (range n))
     :groups (mapv #(.getInt grp (* 4 %)) (range n))
     :delta  (mapv #(.getLong delta (* 8 %)) (range n))}))

